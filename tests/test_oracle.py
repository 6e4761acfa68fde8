"""CPU tests (no GPU): the C++ oracle against (a) the independent numpy restatement written from the
reference sources and (b) the committed golden vectors; plus known-answer tests of the scalar helpers.

The reference has no tests or golden vectors for this path and cannot be compiled here,
so parity is UNPINNED by the reference; these tests are what pins the oracle.
"""
import math
import os

import numpy as np
import pytest

from util import relerr

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(GOLD, name))


# ------------------------------------------------------------------ scalar helpers (KATs)
def test_config_defaults_equal_config_h(orc):
    c = orc.default_config()
    f32 = np.float32
    assert f32(c.pi) == f32(3.141592654) and float(c.pi) == 3.1415927410125732  # config.h:39
    assert c.min_dist_thresh == 0.5 and c.min_landmark_occ == 10 and c.max_state_dim == 30
    assert f32(c.ukf_std_a) == f32(0.2) and f32(c.ukf_std_yaw) == f32(0.2)
    assert f32(c.ukf_kp_robot_pose) == f32(0.001) and c.ukf_kp_landmark_pose == 1.0
    assert f32(c.ukf_kr) == f32(0.2) and f32(c.ukf_kq) == f32(0.001)
    assert f32(c.ekf_kp_robot_pose) == f32(0.001) and c.ekf_kp_landmark_pose == 10000.0
    assert f32(c.ekf_kr) == f32(0.2) and f32(c.ekf_kq) == f32(0.001)
    assert float(f32(c.ekf_kr)) == 0.20000000298023224  # R diagonal as the reference sees it


def test_normalize_angle_kats():
    from oracle import np_restatement as R
    f32 = np.float32
    pi = float(R.PI)
    assert R.normalize_angle(0.0) == 0
    assert R.normalize_angle(1.0) == f32(1.0)
    assert abs(float(R.normalize_angle(4.0)) - (4.0 - 2 * pi)) < 1e-6
    assert abs(float(R.normalize_angle(-4.0)) - (-4.0 + 2 * pi)) < 1e-6
    assert abs(float(R.normalize_angle(7.0)) - (7.0 - 2 * pi)) < 1e-6
    # PI itself is kept (strict comparisons, tools.h:47-48)
    assert R.normalize_angle(R.PI) == R.PI and R.normalize_angle(-R.PI) == -R.PI


def test_first_step_initialises_state_from_measurement(orc):
    """ekf.cpp:87-91: X = Z on the first processed odometry; before any observation message the step is gated."""
    o = orc.Filter("ekf")
    assert not o.step(1.0, 2.0, 0.3, 0.0, 0.0, 1.0)
    o.set_observations([], [])
    assert o.step(1.0, 2.0, 0.3, 0.0, 0.0, 1.0)
    X, Z, P = o.state()
    assert o.N == 3 and Z[0] == 1.0 and Z[1] == 2.0 and Z[2] == float(np.float32(0.3))
    # with X = Z the innovation is zero: X stays at Z
    assert relerr(X, Z) < 1e-15


def test_promotion_on_tenth_sighting(orc):
    o = orc.Filter("ekf")
    for t in range(12):
        o.set_observations([2.0], [0.0])
        o.step(0.0, 0.0, 0.0, 0.0, 0.0, 1.0)
        assert o.N == (3 if t < 9 else 5), t
    X, Z, P = o.state()
    assert abs(X[3] - 2.0) < 1e-3 and abs(X[4]) < 1e-3


# ------------------------------------------------------------------ oracle vs numpy restatement
def _drive(o, r, g, T):
    for t in range(T):
        o.set_observations(g["range"][t], g["bearing"][t])
        r.set_observations(g["range"][t], g["bearing"][t])
        a = (float(g["x"][t]), float(g["y"][t]), g["yaw"][t], float(g["vx"][t]), float(g["az"][t]), g["dt"][t])
        o.step(*a)
        r.step(*a)
        yield t


def test_ekf_oracle_matches_numpy_restatement(orc):
    from oracle import np_restatement as R
    g = _load("cfg1_ekf.npz")
    o, r = orc.Filter("ekf"), R.EkfNp()
    worst = 0.0
    for t in _drive(o, r, g, 120):
        X, Z, P = o.state()
        assert o.N == r.N
        assert list(o.assoc()) == r.assoc, t
        ex, ep = relerr(r.X, X), relerr(r.P, P)
        worst = max(worst, ex, ep)
        assert ex < 1e-10 and ep < 1e-10, (t, ex, ep)
        assert relerr(r.Z, Z) < 1e-15
    assert o.N == 19
    print("EKF oracle vs numpy, 120 messages: worst relerr %.2e" % worst)


def test_ukf_oracle_matches_numpy_restatement(orc):
    """Checked step by step with the numpy filter re-seeded from the oracle every message: the reference UKF is
    chaotic once P has gone indefinite, so free-running agreement of two different summation
    orders is not a meaningful test."""
    from oracle import np_restatement as R
    g = _load("cfg2_ukf.npz")
    o, r = orc.Filter("ukf"), R.UkfNp()
    worst, fails = 0.0, 0
    for t in _drive(o, r, g, 225):
        X, Z, P = o.state()
        assert o.N == r.N and list(o.assoc()) == r.assoc, t
        assert o.llt_fail() == r.llt_fail, t
        fails += o.llt_fail() >= 0
        ex, ep = relerr(r.X, X), relerr(r.P, P)
        worst = max(worst, ex, ep)
        assert ex < 1e-9 and ep < 1e-9, (t, ex, ep)
        r.X, r.P, r.Z = X.copy(), P.copy(), Z.copy()
    assert fails == 3  # messages 101, 209, 219 of the pinned replay hit a non-positive pivot
    print("UKF oracle vs numpy, 225 messages: worst one-step relerr %.2e, LLT failures %d" % (worst, fails))


def test_eigen_llt_rule_small_and_blocked():
    """The LLT restatement: success equals numpy's Cholesky; failure leaves the columns from the failing pivot on
    untouched (unblocked regime, size < 32)."""
    from oracle import np_restatement as R
    rng = np.random.default_rng(0)
    for n in (5, 31, 32, 75, 200):
        G = rng.standard_normal((n, n))
        A = G @ G.T + n * np.eye(n)
        L, k = R.eigen_llt_lower(A)
        assert k == -1 and relerr(L, np.linalg.cholesky(A)) < 1e-12
    A = np.diag([4.0, 1.0, -1.0, 2.0])
    A[3, 0] = A[0, 3] = 0.5
    A[3, 2] = A[2, 3] = 0.25
    L, k = R.eigen_llt_lower(A)
    assert k == 2
    assert L[0, 0] == 2.0 and L[3, 0] == 0.25 and L[1, 1] == 1.0
    assert L[2, 2] == -1.0 and L[3, 2] == 0.25 and L[3, 3] == 2.0  # untouched input values


def test_oracle_llt_failure_rule_matches_restatement(orc):
    g = _load("cfg2_ukf.npz")
    fails = g["llt_fail"]
    assert np.sum(fails >= 0) >= 1, "the pinned replay must exercise the LLT failure path"


# ------------------------------------------------------------------ oracle vs committed golden vectors
@pytest.mark.parametrize("kind,name", [("ekf", "cfg1_ekf.npz"), ("ukf", "cfg2_ukf.npz")])
def test_oracle_reproduces_golden_trace(orc, kind, name):
    g = _load(name)
    steps = list(g["steps"])
    o = orc.Filter(kind)
    T = int(steps[-1]) + 1
    k = 0
    for t in range(T):
        o.set_observations(g["range"][t], g["bearing"][t])
        o.step(float(g["x"][t]), float(g["y"][t]), g["yaw"][t], float(g["vx"][t]), float(g["az"][t]), g["dt"][t])
        if kind == "ukf":
            assert o.llt_fail() == int(g["llt_fail"][t]), t
        if t == steps[k]:
            X, Z, P = o.state()
            n = o.N
            assert n == int(g["N"][k])
            assert np.array_equal(o.assoc(), g["assoc"][k])
            assert relerr(X, g["X"][k][:n]) < 1e-12 and relerr(P, g["P"][k][:n, :n]) < 1e-12, t
            k += 1
    assert k == len(steps)


def test_oracle_batch_runner_reproduces_cfg3_sample(orc):
    g = _load("cfg3_ekf_sample.npz")
    cfg = orc.default_config(max_state_dim=36)
    B, T, L = g["x"].shape[1], g["x"].shape[0], g["range"].shape[2]
    for nthreads in (1, 2):
        sec, X, P, N = orc.run_batch("ekf", cfg, B, T, L, g["x"], g["y"], g["yaw"], g["vx"], g["az"], g["dt"],
                                     g["range"], g["bearing"], nthreads, 36)
        assert np.array_equal(N, g["N"]) and np.all(N == 35)
        assert relerr(X, g["X"]) < 1e-12 and relerr(P, g["P"]) < 1e-12


def test_large_rowsubset_step_equals_dense_slam(orc):
    """with EVERY landmark observed the row-subset joint update is ekf.cpp:300-310 verbatim.
    6 x 6 landmarks here would exceed ASLAM_MAX_OBS; use 3 x 3 (n = 21) with all 9 observed."""
    import awesomeslam_b200 as A
    tx, ty, X0, G, d = A.large_map_state(3)
    P0 = A.dense_P(G, d)
    rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=9)
    n = X0.size
    odom = (0.01, -0.01, 0.002, 0.05, 0.1, 1.0)
    X, P = X0.copy(), P0.copy()
    q, assoc, a00, a10 = orc.ekf_large_step(orc.default_config(), X, P, 1.0, 0.0, *odom, rng, brg)
    assert q == n and sorted(assoc) == list(range(0, 2 * 9, 2))
    # the verbatim dense filter from the same state: Z persistent = h(X0) except what is observed (all of it)
    o = orc.Filter("ekf", max_state_dim=64)
    from oracle import np_restatement as R
    o.load_state(X0, R.measurement(n, X0), P0)
    o.set_observations(rng, brg)
    o.step(*odom)
    Xd, Zd, Pd = o.state()
    assert relerr(X, Xd) < 1e-12 and relerr(P.T, Pd) < 1e-12


def test_large_golden(orc):
    g = _load("large_ekf_l6.npz")
    X, P = g["X0"].copy(), g["P0"].copy()
    q, assoc, a00, a10 = orc.ekf_large_step(orc.default_config(), X, P, 1.0, 0.0, *[float(v) for v in g["odom"]],
                                            g["range"], g["bearing"])
    assert q == int(g["q"]) and np.array_equal(assoc, g["assoc"])
    assert relerr(X, g["X"]) < 1e-13 and relerr(P, g["P_colmajor"]) < 1e-13


def test_ukf_weights_are_float_values(orc):
    from oracle import np_restatement as R
    u = R.UkfNp()
    u._weights(19)
    assert u.w[0] == -6.0 and u.w[1] == float(np.float32(0.5 / 3.0))  # lambda = 3 - 21 = -18
    u._weights(17)
    assert u.w[0] == float(np.float32(-16.0) / np.float32(3.0)) and math.isclose(u.w[0], -5.3333335, rel_tol=1e-7)


def test_vectorised_large_ukf_step_matches_oracle(orc):
    """oracle/np_restatement.ukf_large_step (the full-size cfg5 checker of the GPU tests) against the C++ oracle's
    verbatim UKFSlam::slam at n = 35 (unblocked LLT regime) and n = 131 (blocked)."""
    import awesomeslam_b200 as A
    from oracle import np_restatement as R
    for side, pitch in ((4, 3.0), (8, 2.0)):
        tx, ty, X0, G, d = A.large_map_state(side, pitch=pitch)
        P0 = A.dense_P(G, d)
        n = X0.size
        rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
        Z0 = R.measurement(n, X0)
        odom = (0.01, -0.01, 0.002, 0.05, 0.1, 1.0)
        # one message from a positive-definite state: a second message already finds P - K S K' indefinite (w0 < 0,
        # SURVEY.md section 3.2), where the half-factored L of Eigen's blocked LLT is outside the pinned contract
        for seed in range(3):
            rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8, seed=seed)
            o = orc.Filter("ukf", max_state_dim=100000)
            o.load_state(X0, Z0, P0)
            X, Z, P, assoc, fail = R.ukf_large_step(X0, Z0, P0, *odom, rng, brg)
            o.set_observations(rng, brg)
            o.step(*odom)
            Xo, Zo, Po = o.state()
            assert np.array_equal(assoc, o.assoc()) and fail == o.llt_fail() == -1
            assert relerr(X, Xo) < 1e-11 and relerr(P, Po) < 1e-11 and relerr(Z, Zo) < 1e-15, (n, seed)
