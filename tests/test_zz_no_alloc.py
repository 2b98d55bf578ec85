"""nvllm_op_gemv_noalloc (include/nvllm.h, csrc/noalloc.cu): the allocation-free variant of the
split-K GEMV, its split-K slices in a static device array.  Checked through the C ABI against
the existing entry point for the same kernel (nvllm_op_matmul, kernel = 1), on calls in a row
that share the scratch array.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

from helpers import C, nvllm_pkg, rel_err

L = nvllm_pkg.sub("lib")
TOL = {C.PREC_FP32: 1e-4, C.PREC_BF16: 2e-2}  # the gates of tests/test_gpu_parity.py
MAX_FLOATS = 1 << 20  # NVLLM_NOALLOC_MAX_SCRATCH_FLOATS


def _noalloc(torch, prec, A, W_in_out, stream=None):
    """A [M, K], W [K, N] fp32 numpy -> device tensors the CALLER owns -> C [M, N] numpy."""
    dt = torch.bfloat16 if prec == C.PREC_BF16 else torch.float32
    dA = torch.from_numpy(A).cuda().to(dt).contiguous()
    dW = torch.from_numpy(np.ascontiguousarray(W_in_out.T)).cuda().to(dt).contiguous()  # [out, in]
    dC = torch.full((A.shape[0], W_in_out.shape[1]), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = L.load().nvllm_op_gemv_noalloc(0, prec, A.shape[0], W_in_out.shape[1], A.shape[1],
                                        ctypes.c_void_p(dA.data_ptr()), ctypes.c_void_p(dW.data_ptr()),
                                        ctypes.c_void_p(dC.data_ptr()), ctypes.c_void_p(stream or 0))
    L.check(rc)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return dC.cpu().numpy(), free0 - free1


@pytest.mark.gpu
@pytest.mark.parametrize("prec", [C.PREC_FP32, C.PREC_BF16])
def test_gemv_noalloc_matches_op_matmul_on_calls_in_a_row(prec):
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    lib = L.load()
    # set-up only call (looks the scratch array's device address up), then a warm-up launch so
    # that lazy module loading is not mistaken for an allocation below
    L.check(lib.nvllm_op_gemv_noalloc(0, prec, 0, 0, 0, None, None, None, None))
    rng = np.random.default_rng(7)
    shapes = [(1, 50257, 768),   # GPT-2 LM head, one sequence (cfg1): one slice
              (8, 2304, 768),    # QKV of 8 rows
              (3, 768, 3072),    # FFN down projection: several split-K slices
              (1, 40, 64)]       # tiny, after the large ones: stale scratch must not leak in
    _noalloc(torch, prec, rng.standard_normal((2, 64), dtype=np.float32),
             rng.standard_normal((64, 96), dtype=np.float32))
    for (M, N, K) in shapes:
        A = rng.standard_normal((M, K), dtype=np.float32)
        W = rng.standard_normal((K, N), dtype=np.float32) / np.float32(np.sqrt(K))
        got, _ = _noalloc(torch, prec, A, W)
        ref, _ = L.op_matmul(A, W, None, prec=prec, kernel=1)
        assert np.isfinite(got).all()
        err = rel_err(got, ref)
        assert err <= TOL[prec], f"M={M} N={N} K={K}: rel err {err}"
        # the same kernel, plan and summation order: in fact bit-identical
        assert np.array_equal(got, ref), f"M={M} N={N} K={K}: differs from nvllm_op_matmul (rel {err})"
        # the same call again, straight after: the shared scratch keeps nothing a call depends on
        # (free device memory is compared across THIS call: the first launch of a kernel
        # instantiation may make the driver load its code, which is not an allocation of ours)
        again, grew = _noalloc(torch, prec, A, W)
        assert np.array_equal(again, got)
        assert grew == 0, f"device memory in use grew by {grew} bytes across the call"


@pytest.mark.gpu
def test_gemv_noalloc_rejects_what_does_not_fit():
    torch = pytest.importorskip("torch")
    lib = L.load()
    d = torch.zeros(64, dtype=torch.float32, device="cuda")
    p = ctypes.c_void_p(d.data_ptr())
    # 8 rows x 262144 columns = 2 Mi floats for ONE slice: beyond the 1 Mi-float scratch.  The
    # shape check comes before any launch, so the small buffers are never touched.
    rc = lib.nvllm_op_gemv_noalloc(0, C.PREC_FP32, 8, 262144, 64, p, p, p, None)
    assert rc != 0
    msg = L.last_error()
    assert "scratch" in msg and str(MAX_FLOATS) in msg, msg
    for bad in [(9, 16, 64), (1, 16, 60), (0, 16, 64)]:   # M > 8, K % 8, M < 1
        assert lib.nvllm_op_gemv_noalloc(0, C.PREC_FP32, *bad, p, p, p, None) != 0
    assert lib.nvllm_op_gemv_noalloc(0, C.PREC_FP32, 1, 16, 64, None, p, p, None) != 0
    assert lib.nvllm_op_gemv_noalloc(0, 7, 1, 16, 64, p, p, p, None) != 0
    # and a good call still works afterwards
    A = np.ones((1, 64), dtype=np.float32)
    W = np.ones((64, 16), dtype=np.float32)
    got, _ = _noalloc(torch, C.PREC_FP32, A, W)
    assert np.array_equal(got, np.full((1, 16), 64.0, dtype=np.float32))
