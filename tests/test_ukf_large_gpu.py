"""Parity of the large-map UKF path (cfg5 of BASELINE.json) and of its FP64 dense building blocks, through the C ABI.

 * dense blocks (aslam_dense_*): TMA + DMMA GEMM, blocked Cholesky, triangular solve against numpy, FP64 1e-12
 * one UKF message at n = 131 (64 landmarks, Eigen's blocked-LLT regime) against the C++ oracle's verbatim
   UKFSlam::slam; at n = 35 against the same oracle (unblocked regime)
 * cfg5 full size (n = 2051, 4107 sigma points) against the vectorised numpy restatement, which tests/test_oracle.py
   pins to the oracle, AND against the committed output of the C++ oracle's verbatim UKFSlam::slam at that size
   (tests/golden/cfg5_ukf_large.npz); plus properties: P symmetric, positive definite, status clean
Tolerance: norm-wise relative 1e-9 on X and P (FP64 path); association indices equal.
"""
import ctypes as C

import numpy as np
import pytest

from util import TOL_F64, relerr

pytestmark = pytest.mark.gpu
ODOM = (0.01, -0.01, 0.002, 0.05, 0.1, 1.0)


def _torch():
    import torch
    return torch


def _gemm(aslam, A, B, Cm, alpha=1.0, beta=0.0, lower=False):
    torch = _torch()
    lib = aslam.capi.lib()
    M, K = A.shape
    N = B.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    aslam.capi.check(lib.aslam_dense_gemm_nt(st, M, N, K, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0),
                                             Cm.data_ptr(), Cm.stride(0), alpha, beta, int(lower)))
    torch.cuda.synchronize()


def _padded(torch, rows, cols, rng):
    """rows x cols view of a buffer whose pitch is a multiple of 16 doubles (as the filters allocate)."""
    ld = (cols + 15) // 16 * 16
    buf = torch.zeros((rows, ld), dtype=torch.float64, device="cuda")
    buf[:, :cols] = torch.from_numpy(rng.standard_normal((rows, cols))).cuda()
    return buf[:, :cols]


@pytest.mark.parametrize("M,N,K", [(128, 64, 16), (131, 70, 37), (300, 300, 64), (515, 259, 1030), (2051, 2051, 600)])
def test_dense_gemm_nt(aslam, M, N, K):
    torch = _torch()
    rng = np.random.default_rng(M + N + K)
    A, B, Cm = _padded(torch, M, K, rng), _padded(torch, N, K, rng), _padded(torch, M, N, rng)
    C0 = Cm.clone()
    _gemm(aslam, A, B, Cm, alpha=0.75, beta=-0.5)
    ref = 0.75 * (A @ B.T) - 0.5 * C0
    assert relerr(Cm.cpu().numpy(), ref.cpu().numpy()) < 1e-13


@pytest.mark.parametrize("n,K", [(700, 150), (2051, 520), (1283, 70)])
def test_dense_gemm_lower_only(aslam, n, K):
    torch = _torch()
    rng = np.random.default_rng(3)
    A = _padded(torch, n, K, rng)
    Cm = _padded(torch, n, n, rng)
    C0 = Cm.clone()
    _gemm(aslam, A, A, Cm, alpha=-1.0, beta=1.0, lower=True)
    ref = (C0 - A @ A.T).cpu().numpy()
    got = Cm.cpu().numpy()
    il = np.tril_indices(n)
    assert relerr(got[il], ref[il]) < 1e-13
    # tiles entirely above the diagonal are untouched (128 x 128 or 128 x 64 tiles)
    assert np.array_equal(got[:128, 640:], C0.cpu().numpy()[:128, 640:])


@pytest.mark.parametrize("n", [40, 64, 131, 1000, 2053])
def test_dense_cholesky_and_trsm(aslam, n):
    torch = _torch()
    lib = aslam.capi.lib()
    rng = np.random.default_rng(n)
    G = rng.standard_normal((n, 8))
    P = 0.01 * np.eye(n) + 0.02 * 0.02 * (G @ G.T) + 0.002 * np.diag(rng.random(n))
    ld = (n + 15) // 16 * 16
    Abuf = torch.zeros((n, ld), dtype=torch.float64, device="cuda")
    Abuf[:, :n] = torch.from_numpy(P).cuda()
    nb = (n + 63) // 64
    inv = torch.zeros((nb, 64 * 64), dtype=torch.float64, device="cuda")
    fail = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    aslam.capi.check(lib.aslam_dense_cholesky(st, Abuf.data_ptr(), ld, n, inv.data_ptr(), fail.data_ptr()))
    torch.cuda.synchronize()
    assert int(fail.item()) == -1
    L = np.tril(Abuf[:, :n].cpu().numpy())
    Lref = np.linalg.cholesky(P)
    assert relerr(L, Lref) < 1e-12
    # B <- B L^-T
    m = 200
    B = _padded(torch, m, n, rng)
    B0 = B.cpu().numpy().copy()
    aslam.capi.check(lib.aslam_dense_trsm(st, B.data_ptr(), B.stride(0), m, Abuf.data_ptr(), ld, n, inv.data_ptr()))
    torch.cuda.synchronize()
    ref = np.linalg.solve(Lref, B0.T).T
    assert relerr(B.cpu().numpy(), ref) < 1e-11


def test_dense_cholesky_reports_first_bad_pivot(aslam):
    torch = _torch()
    lib = aslam.capi.lib()
    n = 150
    P = np.eye(n)
    P[100, 100] = -1.0
    Abuf = torch.zeros((n, 160), dtype=torch.float64, device="cuda")
    Abuf[:, :n] = torch.from_numpy(P).cuda()
    inv = torch.zeros((3, 4096), dtype=torch.float64, device="cuda")
    fail = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    aslam.capi.check(lib.aslam_dense_cholesky(torch.cuda.current_stream().cuda_stream, Abuf.data_ptr(), 160, n,
                                              inv.data_ptr(), fail.data_ptr()))
    torch.cuda.synchronize()
    assert int(fail.item()) == 100


def _large_case(aslam, side, pitch, m=8):
    from oracle import np_restatement as R
    tx, ty, X0, G, d = aslam.large_map_state(side, pitch=pitch)
    P0 = aslam.dense_P(G, d)
    rng, brg, idx = aslam.large_observations(tx, ty, (0.0, 0.0, 0.0), m=m)
    Z0 = R.measurement(X0.size, X0)
    return X0, Z0, P0, rng, brg


@pytest.mark.parametrize("side,pitch", [(4, 3.0), (8, 2.0)])
def test_large_ukf_step_matches_oracle(aslam, orc, side, pitch):
    """n = 35 (Eigen's unblocked LLT) and n = 131 (blocked LLT) against the oracle's verbatim UKFSlam::slam."""
    X0, Z0, P0, rng, brg = _large_case(aslam, side, pitch)
    n = X0.size
    tx, ty = aslam.large_map_state(side, pitch=pitch)[:2]
    f = aslam.LargeFilter("ukf", n)
    # one message from a positive-definite state, three different observation draws: a second consecutive message
    # already finds P - K S K' indefinite (w0 < 0), which is outside the large-map contract
    for seed in range(3):
        rng, brg, idx = aslam.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8, seed=seed)
        o = orc.Filter("ukf", max_state_dim=100000)
        o.load_state(X0, Z0, P0)
        f.load_state(X0, P0, Z=Z0)
        o.set_observations(rng, brg)
        o.step(*ODOM)
        assoc = f.step(*ODOM, rng, brg)
        Xo, Zo, Po = o.state()
        X, P = f.get_state()
        assert np.array_equal(assoc, o.assoc())
        status, fail = f.status()
        assert status == 0 and fail == -1 == o.llt_fail()
        ex, ep = relerr(X, Xo), relerr(P, Po)
        print("n=%d seed %d relerr X %.2e P %.2e" % (n, seed, ex, ep))
        assert ex <= TOL_F64 and ep <= TOL_F64
        assert np.array_equal(P, P.T)
    f.close()


def test_large_ukf_default_z_is_h_of_x(aslam, orc):
    X0, Z0, P0, rng, brg = _large_case(aslam, 4, 3.0)
    f = aslam.LargeFilter("ukf", X0.size)
    f.load_state(X0, P0)  # Z = NULL -> h(X)
    f.step(*ODOM, rng, brg)
    o = orc.Filter("ukf", max_state_dim=1000)
    o.load_state(X0, Z0, P0)
    o.set_observations(rng, brg)
    o.step(*ODOM)
    Xo, Zo, Po = o.state()
    X, P = f.get_state()
    assert relerr(X, Xo) <= TOL_F64 and relerr(P, Po) <= TOL_F64
    f.close()


def test_cfg5_large_ukf_full_size(aslam):
    """cfg5: 1024 landmarks over 100 m x 100 m, n = 2051, 4107 sigma points, against the vectorised numpy
    restatement (seconds on the host)."""
    from oracle import np_restatement as R
    X0, Z0, P0, rng, brg = _large_case(aslam, 32, 100.0 / 32)
    n = X0.size
    assert n == 2051
    f = aslam.LargeFilter("ukf", n)
    f.load_state(X0, P0, Z=Z0)
    assoc = f.step(*ODOM, rng, brg)
    X, P = f.get_state()
    status, fail = f.status()
    Xr, Zr, Pr, assoc_r, fail_r = R.ukf_large_step(X0, Z0, P0, *ODOM, rng, brg)
    assert np.array_equal(assoc, assoc_r) and fail == fail_r == -1 and status == 0
    ex, ep = relerr(X, Xr), relerr(P, Pr)
    print("cfg5 n=2051 relerr X %.2e P %.2e" % (ex, ep))
    assert ex <= TOL_F64 and ep <= TOL_F64
    assert np.array_equal(P, P.T)
    # ... and against the C++ oracle's VERBATIM UKFSlam::slam run once at this size (tests/golden/make_cfg5_golden.py,
    # minutes on one host core): X, diag(P) and 32 sampled rows of P
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cfg5_ukf_large.npz"))
    assert np.array_equal(assoc, g["assoc"]) and int(g["llt_fail"]) == -1
    pmax = float(g["Pmax"])
    gx = relerr(X, g["X"])
    gd = float(np.max(np.abs(np.diag(P) - g["Pdiag"])) / pmax)
    gr = float(np.max(np.abs(P[g["rows"]] - g["Prows"])) / pmax)
    print("cfg5 n=2051 vs verbatim oracle golden: relerr X %.2e diag(P) %.2e sampled rows of P %.2e" % (gx, gd, gr))
    assert gx <= TOL_F64 and gd <= TOL_F64 and gr <= TOL_F64
    # a second message: finite, and the LLT verdict (P - K S K' may have gone indefinite) agrees with the host
    f.step(*ODOM, rng, brg)
    X2, P2 = f.get_state()
    status2, fail2 = f.status()
    fail_r2 = R.ukf_large_step(Xr, Zr, Pr, *ODOM, rng, brg)[4]
    assert np.all(np.isfinite(X2)) and (fail2 >= 0) == (fail_r2 >= 0)
    f.close()


def test_large_ukf_flags_indefinite_covariance(aslam):
    X0, Z0, P0, rng, brg = _large_case(aslam, 4, 3.0)
    P0 = P0.copy()
    P0[10, 10] = -1.0
    f = aslam.LargeFilter("ukf", X0.size)
    f.load_state(X0, P0, Z=Z0)
    f.step(*ODOM, rng, brg)
    status, fail = f.status()
    assert status & aslam.capi.ST_LLT_FAILED and fail == 10
    f.close()
