"""cor_debug_gemm on one shape, a few launches: the command ncu captures of a single GEMM variant are taken from.
usage: gemm_probe.py MODE M N K [reps]   (mode 2: tile shape picked from M and N, 3: 128x32 tiles forced)"""
import ctypes as C, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cor_asv_ann_b200 import _lib

mode, M, N, K = (int(v) for v in sys.argv[1:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
lib = _lib.load()
g = torch.Generator(device="cuda").manual_seed(0)
A = torch.randn(M, K, device="cuda", generator=g)
B = torch.randn(N, K, device="cuda", generator=g) * 0.1
out = torch.empty(M, N, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(reps):
    _lib.check(lib.cor_debug_gemm(mode, M, N, K, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(out.data_ptr()), st))
torch.cuda.synchronize()
print("ok", mode, M, N, K)
