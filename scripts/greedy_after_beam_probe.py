"""Does a beam call before it slow the fast-greedy loop graph down?  (bench.py's default run measured 191 ms per greedy step
after the beam legs against 147 ms alone.)  argv[1] = 1: run the beam calls first."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cor_asv_ann_b200.engine import Engine

beam_first = len(sys.argv) > 1 and sys.argv[1] == "1"
budget_gb = float(sys.argv[2]) if len(sys.argv) > 2 else 72.0
eng = Engine(dict(bench.C2, window_width=5), bench.build_weights(bench.C2, "U"), device=0, eos_index=1)
T = bench.LINE_LEN + 1
if beam_first:
    idx = torch.as_tensor(bench.make_indices(9472, 1), device="cuda")
    row_T = torch.full((9472,), T, dtype=torch.int32, device="cuda")
    bp = eng.beam_params(**bench.BEAM_C2)
    out = None
    for _ in range(4):
        out = eng.beam_device(idx, None, row_T, bp, budget_bytes=int(budget_gb * 2**30), out=out)
    torch.cuda.synchronize()
L = 16384
idx = [torch.as_tensor(bench.make_indices(L, 10 + s), device="cuda") for s in range(4)]
row_T = torch.full((L,), T, dtype=torch.int32, device="cuda")
for mode in ("graph", "host"):
    if mode == "host":
        os.environ["COR_NO_GRAPH"] = "1"
    for s in range(3):
        eng.encode(idx[s % 4], None, row_T, want_outputs=False); eng.greedy()
    torch.cuda.synchronize()
    per = []
    for s in range(5):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(); eng.encode(idx[s % 4], None, row_T, want_outputs=False); e1.record(); eng.greedy(); e2.record()
        torch.cuda.synchronize()
        per.append((round(e0.elapsed_time(e1), 1), round(e1.elapsed_time(e2), 1)))
    print("beam_first=%s ws=%.1f GB %s loop: (encode ms, decode ms) per step: %s" % (beam_first, eng._ws.numel() / 2**30, mode, per))
