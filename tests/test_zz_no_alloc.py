"""aslam_dense_gemm_nt_tf32_static: the TF32 (tcgen05) GEMM with its scratch in a file-scope static __device__ array
instead of cudaMalloc (csrc/no_alloc.cu; no counterpart in the reference).

Called through the C ABI and compared with the existing, allocating entry point aslam_dense_gemm_nt_tf32 on the same
inputs, on two calls in a row (the second smaller and with the other pass count, so it finds the first call's contents
in the shared scratch). Tolerance: the TF32 GEMM tolerances of tests/util.py (the two entry points run the same kernels
in the same order, so the difference is in fact zero -- printed). Plus: a size beyond the static scratch is refused with
a clear error and leaves the output untouched. The file sorts last on purpose.
"""
import numpy as np
import pytest

from util import TOL_TF32_GEMM, TOL_TF32X3_GEMM, relerr

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    return torch


def _operand(torch, rows, cols, rng):
    """rows x cols view of a device buffer whose pitch is a multiple of 16 doubles (as the filters allocate)."""
    ld = (cols + 15) // 16 * 16
    buf = torch.zeros((rows, ld), dtype=torch.float64, device="cuda")
    buf[:, :cols] = torch.from_numpy(rng.standard_normal((rows, cols))).cuda()
    return buf[:, :cols]


def _call(aslam, name, A, B, Cm, passes, alpha, beta, lower=False):
    torch = _torch()
    fn = getattr(aslam.capi.lib(), name)
    M, K = A.shape
    N = B.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    rc = fn(st, M, N, K, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), Cm.data_ptr(), Cm.stride(0), alpha, beta,
            int(lower), passes)
    torch.cuda.synchronize()
    return rc


def test_static_scratch_gemm_equals_the_allocating_entry_point_twice_in_a_row(aslam):
    torch = _torch()
    rng = np.random.default_rng(20261105)
    # (M, N, K, passes): 8.0 of the 8 MiB for the first call, then a smaller, ragged one with the other pass count
    for call, (M, N, K, passes) in enumerate([(515, 509, 1000, 3), (131, 200, 77, 1)]):
        A, B = _operand(torch, M, K, rng), _operand(torch, N, K, rng)
        C0 = _operand(torch, M, N, rng)
        want, got = C0.clone(), C0.clone()
        aslam.capi.check(_call(aslam, "aslam_dense_gemm_nt_tf32", A, B, want, passes, 0.75, -0.5))
        aslam.capi.check(_call(aslam, "aslam_dense_gemm_nt_tf32_static", A, B, got, passes, 0.75, -0.5))
        w, g = want.cpu().numpy(), got.cpu().numpy()
        tol = TOL_TF32_GEMM if passes == 1 else TOL_TF32X3_GEMM
        err = relerr(g, w)
        ref = (0.75 * (A @ B.T) - 0.5 * C0).cpu().numpy()
        print("call %d: %dx%dx%d passes %d: static vs allocating relerr %.2e (bit-equal: %s), vs FP64 %.2e"
              % (call, M, N, K, passes, err, np.array_equal(g, w), relerr(g, ref)))
        assert err <= tol
        assert relerr(g, ref) <= tol


def test_static_scratch_gemm_refuses_sizes_beyond_its_scratch(aslam):
    torch = _torch()
    rng = np.random.default_rng(7)
    M = N = 1024
    K = 1100  # padded to 1120: 2 x 4.6 MB x 2 (hi + lo) > 8 MiB with three passes, and 9.2 MB > 8 MiB with one
    A, B = _operand(torch, M, K, rng), _operand(torch, N, K, rng)
    Cm = torch.full((M, N), 7.0, dtype=torch.float64, device="cuda")
    for passes in (3, 1):
        rc = _call(aslam, "aslam_dense_gemm_nt_tf32_static", A, B, Cm, passes, 1.0, 0.0)
        assert rc == aslam.capi.ASLAM_ERR_UNSUPPORTED
        msg = aslam.capi.lib().aslam_last_error().decode()
        assert "static scratch holds 8388608" in msg and "M = 1024" in msg
    assert bool((Cm == 7.0).all())  # nothing was launched
    # ... and the entry point still works afterwards, at a size that fits
    A, B = _operand(torch, 128, 64, rng), _operand(torch, 128, 64, rng)
    out = torch.zeros((128, 128), dtype=torch.float64, device="cuda")
    aslam.capi.check(_call(aslam, "aslam_dense_gemm_nt_tf32_static", A, B, out, 3, 1.0, 0.0))
    assert relerr(out.cpu().numpy(), (A @ B.T).cpu().numpy()) <= TOL_TF32X3_GEMM
