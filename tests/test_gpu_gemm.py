"""The three GEMM kernels against a float64 matmul, through the C-ABI test hook cor_debug_gemm."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(mode, M, N, K, seed=0):
    from cor_asv_ann_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn(M, K, device="cuda", generator=g)
    B = torch.randn(N, K, device="cuda", generator=g) * 0.1
    Cm = torch.full((M, N), float("nan"), device="cuda")
    _lib.check(lib.cor_debug_gemm(mode, M, N, K, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()),
                                  C.c_void_p(Cm.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    ref = A.double() @ B.double().T
    scale = (A.double().abs() @ B.double().abs().T)
    return ((Cm.double() - ref).abs() / scale).max().item()


@pytest.mark.parametrize("shape", [(128, 128, 16), (300, 256, 64), (77, 512, 512)])
def test_ffma_gemm(shape):
    err = _run(1, *shape)
    print("ffma %s: %.3g" % (shape, err))
    assert err < 1e-6


@pytest.mark.parametrize("shape", [(1, 16, 64), (8, 512, 512), (32, 2048, 1536), (33, 1024, 576), (64, 2048, 1024),
                                   (61, 40, 128), (130, 48, 64), (300, 2048, 512)])
def test_tc_narrow_gemm(shape):
    """The page-sized (small M) 128x32-tile variant of the tcgen05 kernel, with row and column tails."""
    err = _run(3, *shape)
    print("narrow %s: %.3g" % (shape, err))
    assert err < 2e-6


@pytest.mark.parametrize("shape", [(128, 256, 64), (128, 256, 128), (256, 512, 512), (1000, 2048, 1536),
                                   (77, 128, 64), (333, 96, 192), (16384, 512, 512),
                                   # enough 256-row tiles for the CTA-pair kernel (cta_group::2), with M and N tails
                                   (18999, 2048, 1024), (9473, 768, 192), (40000, 96, 64)])
def test_tcgen05_3xfp16_gemm(shape):
    """Relative to sum |a||b| the 3-pass fp16 hi/lo result must be fp32-grade (one fp16/TF32 pass is ~5e-4)."""
    err = _run(2, *shape)
    print("tcgen05 3xFP16 %s: max |err| / sum|a||b| = %.3g" % (shape, err))
    assert err < 1e-5, err


def test_tcgen05_small_magnitudes():
    """fp16 residuals of small operands sit on the subnormal grid: the ABSOLUTE error stays ~1e-7."""
    from cor_asv_ann_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 256, 256, 1024
    A = torch.randn(M, K, device="cuda", generator=g) * 1e-3          # like small softmax / state entries
    B = torch.randn(N, K, device="cuda", generator=g) * 0.05          # like glorot weights at W=512
    Cm = torch.empty(M, N, device="cuda")
    _lib.check(lib.cor_debug_gemm(2, M, N, K, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()),
                                  C.c_void_p(Cm.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    err = (Cm.double() - A.double() @ B.double().T).abs().max().item()
    print("abs err %.3g" % err)
    assert err < 2e-7


def test_tcgen05_small_magnitude_weights():
    """Small WEIGHTS (the B operand: Keras-init embedding sigma 0.001, and the tables derived from it): the residual of
    every element is subnormal, the representation keeps ~2^-15 relative, and the product error is bounded in
    absolute terms -- |err| <= K * max|a| * 2^-25-ish, far below the 1e-4 parity budget -- NOT in fp32-relative ones."""
    from cor_asv_ann_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(6)
    M, N, K = 256, 512, 512
    A = torch.tanh(torch.randn(M, K, device="cuda", generator=g))     # LSTM outputs in (-1, 1)
    B = torch.randn(N, K, device="cuda", generator=g) * 1e-3          # embedding rows, sigma 0.001
    Cm = torch.empty(M, N, device="cuda")
    _lib.check(lib.cor_debug_gemm(2, M, N, K, C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()),
                                  C.c_void_p(Cm.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    ref = A.double() @ B.double().T
    err = (Cm.double() - ref).abs().max().item()
    rel = err / ref.abs().max().item()
    print("abs err %.3g (relative to the largest product %.3g)" % (err, rel))
    assert err < 3e-6          # ~ sqrt(K) * max|a| * 2^-25-ish residual grid of B plus the fp32 accumulation: 30x below the parity budget
    assert rel < 1e-4          # the stated grade for small operands: absolute, i.e. a relative 2^-15..2^-13 here


@pytest.mark.parametrize("shape", [(256, 512, 512), (77, 128, 64), (18999, 2048, 1024), (32, 2048, 1536)])
def test_tcgen05_single_pass_gemm(shape):
    """gemm_mode 3 (fp16 heads only, one tensor pass): the error of ONE rounding of each operand to 11
    significand bits, two to three orders above the three-pass product and well inside 1e-3."""
    err1, err3 = _run(4, *shape), _run(2, *shape)
    print("tcgen05 1xFP16 %s: %.3g (three-pass %.3g)" % (shape, err1, err3))
    assert err3 < 1e-5 < err1 < 1e-3, (err1, err3)
