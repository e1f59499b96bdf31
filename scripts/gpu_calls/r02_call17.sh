# round 2, call 17: where did the beamed page through the class go from 8.3 to 25.7 ms?
mkdir -p gpurun_out
python scripts/page_beam_probe.py 32 10
COR_NO_EPI_TMA=1 python scripts/page_beam_probe.py 32 10
python - <<'PY'
import time, numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
import bench
from cor_asv_ann_b200.synthetic import synthetic_alphabet
mapping = synthetic_alphabet(512)
s = bench.make_s2s(bench.C2, bench.build_weights(bench.C2, "U"), mapping, 0)
i2c = np.array([ord(mapping[1][i]) if mapping[1][i] else 0 for i in range(512)], dtype=np.uint32)
page = [i2c[bench.make_indices(1, 5000 + j)[0, :, 0]].tobytes().decode("utf-32-le") for j in range(32)]
s.rejection_threshold, s.beam_width_in = 0.5, 4
eng = s._get_engine()
for k in range(8):
    eng.launch_count(reset=True)
    t0 = time.perf_counter()
    s.correct_lines(page, None, fast=False, greedy=False)
    dt = time.perf_counter() - t0
    print("call %d: %.2f ms, kernels %d, host launches %d" % (k, dt * 1e3, eng.launch_count(), eng.host_launch_count()))
PY
( timeout 600 python -m pytest tests/test_gpu_loop_graph.py -m gpu -x -q 2>&1 | tail -4 )
