"""cor_gemm_noalloc (csrc/noalloc.cu): the tcgen05 GEMM with its split operands in a file-scope device array instead of
cudaMalloc'ed buffers, against the existing entry point cor_debug_gemm(mode 2), through the C-ABI.

No counterpart in the reference.  Collected last on purpose (`zz`).
Both entry points run the same split arithmetic and the same kernel, so the results are
expected to be IDENTICAL; the asserted tolerance is the one tests/test_gpu_gemm.py uses for this kernel
(|err| <= 1e-5 * sum |a||b|), against the existing entry point and against a float64 product.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5          # relative to sum_k |a_k||b_k|, as in tests/test_gpu_gemm.py::test_tcgen05_3xfp16_gemm


def _operands(M, N, K, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn(M, K, device="cuda", generator=g)
    B = torch.randn(N, K, device="cuda", generator=g) * 0.1
    return A, B


def _call(fn, M, N, K, A, B):
    from cor_asv_ann_b200 import _lib
    out = torch.full((M, N), float("nan"), device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(fn(M, N, K, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(out.data_ptr()), st))
    torch.cuda.synchronize()
    return out


def _check(lib, M, N, K, seed):
    A, B = _operands(M, N, K, seed)
    got = _call(lib.cor_gemm_noalloc, M, N, K, A, B)
    ref = _call(lambda *a: lib.cor_debug_gemm(2, *a), M, N, K, A, B)
    scale = A.double().abs() @ B.double().abs().T
    err_entry = ((got.double() - ref.double()).abs() / scale).max().item()
    err_f64 = ((got.double() - A.double() @ B.double().T).abs() / scale).max().item()
    print("no-alloc %s: vs cor_debug_gemm %.3g (identical: %s), vs float64 %.3g" % ((M, N, K), err_entry, bool((got == ref).all()), err_f64))
    assert torch.isfinite(got).all()
    assert err_entry < TOL and err_f64 < TOL, (err_entry, err_f64)


def test_no_alloc_gemm_two_calls_in_a_row():
    """Two calls in a row on the shared scratch array, the second with a different (smaller) geometry so that it reads
    a scratch that still holds the first call's operands beyond its own."""
    from cor_asv_ann_b200 import _lib
    lib = _lib.load()
    assert lib.cor_gemm_noalloc_max_bytes() == 16 << 20
    _check(lib, 1000, 2048, 1024, seed=11)      # 4 * (M + N) * K = 12.5 MB of the 16 MiB
    _check(lib, 77, 512, 512, seed=12)
    _check(lib, 1000, 2048, 1024, seed=13)      # and the first geometry again, new operands


def test_no_alloc_gemm_rejects_what_does_not_fit():
    from cor_asv_ann_b200 import _lib
    lib = _lib.load()
    M, N, K = 2048, 2048, 1024                   # 16 MiB + nothing to spare: 4 * 4096 * 1024 = 16 MiB fits exactly ...
    A, B = _operands(M, N, K, 1)
    _call(lib.cor_gemm_noalloc, M, N, K, A, B)
    M = 2049                                     # ... one more row does not
    A, B = _operands(M, N, K, 2)
    out = torch.zeros(M, N, device="cuda")
    rc = lib.cor_gemm_noalloc(M, N, K, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(out.data_ptr()),
                              C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.cor_last_error().decode()
    assert "16 MiB" in msg, msg
    torch.cuda.synchronize()
    assert (out == 0).all()                      # nothing was launched
    rc = lib.cor_gemm_noalloc(64, 64, 100, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(out.data_ptr()), None)
    assert rc != 0 and "multiple of 64" in lib.cor_last_error().decode()
