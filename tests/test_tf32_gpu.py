"""The TF32 variant (north_star: "FP64/TF32 tensor-core tiles via TMA"; SURVEY 8d cfg5 "TF32 variant: separate row,
tolerance stated separately"): tcgen05.mma kind::tf32 with the accumulator in tensor memory, csrc/gemm_tf32_tcgen05.cu.

 * the GEMM by itself against numpy FP64, single pass (operands rounded to TF32) and 3-pass hi/lo split
 * one large-map UKF message with the three sigma-point products in TF32 / 3xTF32 against the FP64 path of the same
   library (which the other tests hold to 1e-9 against the oracle), at n = 131, 515 and at cfg5 size n = 2051
Tolerances: util.TOL_TF32 / TOL_TF32X3 (norm-wise relative, the measured values are printed). The FP64 path stays the
parity path; this one is opt-in (aslam_large_ukf_set_precision).
"""
import ctypes as C

import numpy as np
import pytest

from test_ukf_large_gpu import ODOM, _large_case, _padded, _torch
from util import TOL_TF32, TOL_TF32X3, TOL_TF32_GEMM, TOL_TF32X3_GEMM, relerr

pytestmark = pytest.mark.gpu


def _gemm_tf32(aslam, A, B, Cm, passes, alpha=1.0, beta=0.0, lower=False):
    torch = _torch()
    lib = aslam.capi.lib()
    M, K = A.shape
    N = B.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    aslam.capi.check(lib.aslam_dense_gemm_nt_tf32(st, M, N, K, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0),
                                                  Cm.data_ptr(), Cm.stride(0), alpha, beta, int(lower), passes))
    torch.cuda.synchronize()


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (131, 200, 77), (300, 300, 4107), (515, 259, 1030)])
@pytest.mark.parametrize("passes", [1, 3])
def test_tf32_gemm_against_numpy(aslam, M, N, K, passes):
    torch = _torch()
    rng = np.random.default_rng(M + N + K)
    A, B, Cm = _padded(torch, M, K, rng), _padded(torch, N, K, rng), _padded(torch, M, N, rng)
    C0 = Cm.clone()
    _gemm_tf32(aslam, A, B, Cm, passes, alpha=0.75, beta=-0.5)
    ref = 0.75 * (A @ B.T) - 0.5 * C0
    err = relerr(Cm.cpu().numpy(), ref.cpu().numpy())
    print("tf32 gemm %dx%dx%d passes %d: relerr %.2e" % (M, N, K, passes, err))
    assert err <= (TOL_TF32_GEMM if passes == 1 else TOL_TF32X3_GEMM)


def test_tf32_gemm_lower_only_leaves_the_upper_tiles(aslam):
    torch = _torch()
    rng = np.random.default_rng(3)
    n, K = 700, 150
    A = _padded(torch, n, K, rng)
    Cm = torch.full((n, n), 7.0, dtype=torch.float64, device="cuda")
    _gemm_tf32(aslam, A, A, Cm, 3, lower=True)
    ref = (A @ A.T).cpu().numpy()
    got = Cm.cpu().numpy()
    il = np.tril_indices(n)
    assert relerr(got[il], ref[il]) <= TOL_TF32X3_GEMM
    assert np.all(got[np.triu_indices(n, 1)] == 7.0)  # never written above the diagonal


def test_tf32_probe_reports_a_plausible_peak(aslam):
    tf = C.c_double()
    aslam.capi.check(aslam.capi.lib().aslam_probe_tf32_tflops(0, C.byref(tf)))
    print("kind::tf32 peak probe: %.0f TFLOP/s" % tf.value)
    assert 200.0 < tf.value < 2000.0


@pytest.mark.parametrize("side,pitch", [(8, 3.0), (16, 100.0 / 16), (32, 100.0 / 32)])
def test_large_ukf_step_tf32_against_the_fp64_path(aslam, side, pitch):
    X0, Z0, P0, rng, brg = _large_case(aslam, side, pitch)
    n = X0.size
    out = {}
    for mode in (0, 1, 3):
        f = aslam.LargeFilter("ukf", n)
        f.set_precision(mode)
        f.load_state(X0, P0, Z=Z0)
        f.step(*ODOM, rng, brg)
        X, P = f.get_state()
        out[mode] = (X, P, f.status())
        f.close()
    assert out[0][2] == (0, -1)
    for mode, tx, tp in ((1, TOL_TF32[0], TOL_TF32[1]), (3, TOL_TF32X3[0], TOL_TF32X3[1])):
        ex, ep = relerr(out[mode][0], out[0][0]), relerr(out[mode][1], out[0][1])
        print("large UKF n=%d precision %d: relerr X %.2e P %.2e status %s" % (n, mode, ex, ep, out[mode][2]))
        assert out[mode][2] == (0, -1)
        assert ex <= tx and ep <= tp
