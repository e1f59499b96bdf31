"""One device-resident beam call of a bench workload (bench.WORKLOADS) on `lines` lines, `reps` times: the command the
launch lists / ncu captures of the competitive and stress brackets are taken from (no warm-up floor, no other legs)."""
import os, sys, time, json
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cor_asv_ann_b200.engine import Engine

workload = sys.argv[1] if len(sys.argv) > 1 else "beam_competitive"
n_lines = int(sys.argv[2]) if len(sys.argv) > 2 else 256
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
budget = int(float(sys.argv[4]) * (1 << 30)) if len(sys.argv) > 4 else 16 << 30
spec = bench.WORKLOADS[workload]
eng = Engine(dict(spec["topo"], window_width=5), bench.build_weights(spec["topo"], spec["variant"]), device=0, eos_index=1)
T = bench.LINE_LEN + 1
idx = torch.as_tensor(bench.make_indices(n_lines, spec["seed"], spec["topo"]["voc_size"]), device="cuda")
row_T = torch.full((n_lines,), T, dtype=torch.int32, device="cuda")
bp = eng.beam_params(**spec["beam"])
out = None
eng.launch_count(reset=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
each = []
for s in range(reps):
    t1 = time.perf_counter()
    out = eng.beam_device(idx, None, row_T, bp, budget_bytes=budget, out=out)
    torch.cuda.synchronize()
    each.append(round((time.perf_counter() - t1) * 1e3, 1))
dt = (time.perf_counter() - t0) / reps
st = out[5].cpu().numpy()
print(json.dumps({"workload": workload, "lines": n_lines, "ms_per_call": dt * 1e3, "kernels_per_call": eng.launch_count() / reps,
                  "iterations_max": int(st[:, 0].max()), "hyp_steps": int(st[:, 1].sum()), "hyp_steps_per_s": float(st[:, 1].sum() / dt),
                  "fringe_peak_max": int(st[:, 7].max()), "ms_each_call": each}))
