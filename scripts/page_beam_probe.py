"""One OCR-D page (32 lines of 80 characters) through the beamed decoder, a few calls: the command the page-regime
launch lists and ncu captures of the beam search are taken from (bench.py --page-only also runs the cooperative
page decoder, which ncu cannot replay)."""
import os, sys, time, json
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cor_asv_ann_b200.engine import Engine

n_lines = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
eng = Engine(dict(bench.C2, window_width=5), bench.build_weights(bench.C2, "U"), device=0, eos_index=1)
T = bench.LINE_LEN + 1
idx = [torch.as_tensor(bench.make_indices(n_lines, 2000 + s)).pin_memory() for s in range(4)]
host_T = torch.full((n_lines,), T, dtype=torch.int32).pin_memory()
bp = eng.beam_params(**bench.BEAM_C2)
out = None
for s in range(2):
    out = eng.correct_beam_host(idx[s % 4], None, host_T, bp, budget_bytes=2 << 30, out=out)
torch.cuda.synchronize()
eng.launch_count(reset=True)
t0 = time.perf_counter()
for s in range(reps):
    out = eng.correct_beam_host(idx[s % 4], None, host_T, bp, budget_bytes=2 << 30, out=out)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / reps
print(json.dumps({"lines": n_lines, "ms_per_call": dt * 1e3, "kernels_per_call": eng.launch_count() / reps,
                  "iterations": int(out[5][:, 0].max())}))
