"""The CUDA path (through the C ABI) against the COMMITTED golden vectors of tests/golden/ -- no oracle and no
/root/reference at run time. FP64 tolerance 1e-9 (norm-wise, X and P separately); N, association and the LLT
failure column must be equal."""
import os

import numpy as np
import pytest

from util import TOL_F64, relerr

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("kind,name", [("ekf", "cfg1_ekf.npz"), ("ukf", "cfg2_ukf.npz")])
def test_single_filter_golden_trace(aslam, kind, name):
    g = np.load(os.path.join(GOLD, name))
    steps = [int(s) for s in g["steps"]]
    if kind == "ukf":
        # The reference UKF is chaotic once P has gone indefinite and the trace holds no Z to
        # re-seed from, so the free-running comparison stops at message 12; the 1000-message UKF parity with counted
        # re-synchronisations against the live oracle is tests/test_ukf_batched_gpu.py.
        steps = [s for s in steps if s <= 12]
    f = aslam.BatchedFilter(kind, 1, 0)
    k = 0
    for t in range(steps[-1] + 1):
        f.set_observations(g["range"][t], g["bearing"][t])
        f.step(g["x"][t:t + 1], g["y"][t:t + 1], g["yaw"][t:t + 1], g["vx"][t:t + 1], g["az"][t:t + 1], g["dt"][t:t + 1])
        if kind == "ukf":
            assert int(f.get_llt_fail()[0]) == int(g["llt_fail"][t]), t
        if t == steps[k]:
            s = f.get_state()
            n = int(g["N"][k])
            assert int(s["N"][0]) == n
            assert np.array_equal(s["assoc"], g["assoc"][k][:s["assoc"].size])
            ex, ep = relerr(s["X"][0, :n], g["X"][k][:n]), relerr(s["P"][0, :n, :n], g["P"][k][:n, :n])
            assert ex <= TOL_F64 and ep <= TOL_F64, (t, ex, ep)
            k += 1
    assert k == len(steps)
    f.close()


def test_cfg3_sample_golden(aslam):
    g = np.load(os.path.join(GOLD, "cfg3_ekf_sample.npz"))
    T, B, L = g["range"].shape
    f = aslam.BatchedFilter("ekf", B, 0, max_state_dim=36)
    f.run_host(T, g["x"], g["y"], g["yaw"], g["vx"], g["az"], g["dt"], g["range"], g["bearing"], L)
    s = f.get_state()
    assert np.array_equal(s["N"], g["N"])
    assert relerr(s["X"], g["X"]) <= TOL_F64 and relerr(s["P"], g["P"]) <= TOL_F64
    f.close()


def test_large_ekf_golden(aslam):
    g = np.load(os.path.join(GOLD, "large_ekf_l6.npz"))
    f = aslam.LargeFilter("ekf", g["X0"].size)
    f.load_state(g["X0"], g["P0"])
    assoc = f.step(*[float(v) for v in g["odom"]], g["range"], g["bearing"])
    X, P = f.get_state()
    assert np.array_equal(assoc, g["assoc"])
    assert relerr(X, g["X"]) <= TOL_F64 and relerr(P, g["P_colmajor"].T) <= TOL_F64
    f.close()
