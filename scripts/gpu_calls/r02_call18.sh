# round 2, call 18: beamed page through the class AFTER a predict() run (bench order)
mkdir -p gpurun_out
python - <<'PY'
import time, numpy as np, torch, sys, os, tempfile
sys.path.insert(0, os.getcwd())
import bench
from cor_asv_ann_b200.synthetic import synthetic_alphabet
mapping = synthetic_alphabet(512)
s = bench.make_s2s(bench.C2, bench.build_weights(bench.C2, "U"), mapping, 0)
i2c = np.array([ord(mapping[1][i]) if mapping[1][i] else 0 for i in range(512)], dtype=np.uint32)
with tempfile.TemporaryDirectory() as tmp:
    path = os.path.join(tmp, "l.txt")
    idx = bench.make_indices(8192, 3)[:, :, 0]
    open(path, "w", encoding="utf-8").write(i2c[idx].tobytes().decode("utf-32-le"))
    s.superbatch = 16
    t0 = time.perf_counter(); n = sum(len(l) for _, l, _ in s.predict([path], fast=True, greedy=True)); print("predict", n, time.perf_counter() - t0)
page = [i2c[bench.make_indices(1, 5000 + j)[0, :, 0]].tobytes().decode("utf-32-le") for j in range(32)]
s.rejection_threshold, s.beam_width_in = 0.5, 4
eng = s._get_engine()
for fast in (True, False):
    for k in range(6):
        eng.launch_count(reset=True)
        t0 = time.perf_counter()
        s.correct_lines(page, None, fast=fast, greedy=fast)
        dt = time.perf_counter() - t0
        print("fast %s call %d: %.2f ms, kernels %d, host launches %d, ws %.1f GB" % (fast, k, dt * 1e3, eng.launch_count(), eng.host_launch_count(), eng._ws.numel() / 2**30))
PY
