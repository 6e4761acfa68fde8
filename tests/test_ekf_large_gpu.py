"""Parity of the large-map EKF (row-subset joint update, streaming covariance kernel) through the
C ABI against the CPU oracle. Tolerance: norm-wise relative error <= 1e-9 on X and P (FP64);
association indices EQUAL."""
import numpy as np
import pytest

from util import TOL_F64, relerr

pytestmark = pytest.mark.gpu


def _trajectory(steps):
    """A few odometry messages: stationary first, then the cfg arc (v=0.05, w=0.1, dt=1)."""
    x = y = th = 0.0
    out = []
    for t in range(steps):
        v, w = (0.0, 0.0) if t < 1 else (0.05, 0.1)
        if w:
            x += v / w * (np.sin(th + w) - np.sin(th))
            y -= v / w * (np.cos(th + w) - np.cos(th))
            th += w
        out.append((x, y, th, v, w))
    return out


def _run_large(aslam, orc, L_side, m, steps, seed=4):
    tx, ty, X0, G, d = aslam.large_map_state(L_side, seed=seed)
    P0 = aslam.dense_P(G, d)  # exactly symmetric: row-major == column-major
    n = X0.size
    f = aslam.LargeFilter("ekf", n)
    f.load_state_colmajor(X0, P0)
    Xo, Po = X0.copy(), P0.copy()
    cfg = orc.default_config()
    a00, a10 = 1.0, 0.0
    worst = [0.0, 0.0]
    rs = np.random.default_rng(11)
    for t, (px, py, pth, v, w) in enumerate(_trajectory(steps)):
        rng, brg, idx = aslam.large_observations(tx, ty, (px, py, pth), m=m, seed=100 + t)
        ox, oy = px + 0.01 * rs.standard_normal(), py + 0.01 * rs.standard_normal()
        yaw = np.float32(pth + 0.01 * rs.standard_normal())
        assoc = f.step(ox, oy, yaw, v, w, 1.0, rng, brg)
        q, assoc_o, a00, a10 = orc.ekf_large_step(cfg, Xo, Po, a00, a10, ox, oy, yaw, v, w, 1.0, rng, brg)
        assert np.array_equal(assoc, assoc_o), "association differs at step %d" % t
        assert np.array_equal(assoc // 2, idx), "synthetic observations should hit their own landmarks"
        X, P = f.get_state()
        ex, ep = relerr(X, Xo), relerr(P, Po.T)  # Po is column-major: Po[j, i] = P[i, j]
        worst = [max(worst[0], ex), max(worst[1], ep)]
        assert ex <= TOL_F64 and ep <= TOL_F64, "step %d: relerr X %.3e P %.3e" % (t, ex, ep)
    f.close()
    return worst


def test_large_ekf_small_map_q19(aslam, orc):
    """n = 131 (8x8 landmarks), 8 observed per step -> q = 19: the register-tiled streaming kernel."""
    print("worst", _run_large(aslam, orc, 8, 8, 6))


def test_large_ekf_ragged_sizes(aslam, orc):
    """n not a multiple of anything: 5x5 (n=53), 13x13 (n=341), 23x23 (n=1061); m = 0, 1, 3, 17, 40."""
    print(_run_large(aslam, orc, 5, 3, 3))
    print(_run_large(aslam, orc, 13, 17, 3))   # q = 37 -> the 1-row x 40 variant
    print(_run_large(aslam, orc, 23, 40, 2))   # q = 83 -> generic variant
    print(_run_large(aslam, orc, 13, 1, 2))
    print(_run_large(aslam, orc, 5, 0, 2))     # no observation: odometry-only update, q = 3


def test_large_ekf_all_observed_equals_verbatim_dense_slam(aslam, orc):
    """With every landmark observed the row-subset update IS ekf.cpp:300-310: compare with the
    oracle's verbatim dense EKFSlam::slam (7x7 landmarks, n = 101, q = 101)."""
    tx, ty, X0, G, d = aslam.large_map_state(7, seed=5)
    P0 = aslam.dense_P(G, d)
    n = X0.size
    f = aslam.LargeFilter("ekf", n)
    f.load_state_colmajor(X0, P0)
    o = orc.Filter("ekf", max_state_dim=n + 1)
    o.load_state(X0, X0.copy(), P0)
    for t, (px, py, pth, v, w) in enumerate(_trajectory(4)):
        rng, brg, idx = aslam.large_observations(tx, ty, (px, py, pth), m=49, seed=7 + t)
        assoc = f.step(px, py, np.float32(pth), v, w, 1.0, rng, brg)
        o.set_observations(rng, brg)
        o.step(px, py, np.float32(pth), v, w, 1.0)
        assert np.array_equal(assoc, o.assoc())
        X, P = f.get_state()
        Xo, Zo, Po = o.state()
        assert o.N == n
        ex, ep = relerr(X, Xo), relerr(P, Po)
        assert ex <= TOL_F64 and ep <= TOL_F64, "step %d: relerr X %.3e P %.3e" % (t, ex, ep)
    f.close()


class _DevArray:
    """CUDA array interface view of a device pointer (so torch can wrap library-owned memory)."""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 2}


def test_large_ekf_row_sharded_two_handles_one_gpu(aslam, orc):
    """Row-block sharding emulated on ONE GPU: two handles own rows [0,k) and
    [k,n); their exchange buffers are summed (the all-reduce) and phase 2 runs on each. Every
    exchanged element has one owner, so the result is BIT-IDENTICAL to the unsharded handle."""
    import torch
    tx, ty, X0, G, d = aslam.large_map_state(9, seed=6)
    P0 = aslam.dense_P(G, d)
    n = X0.size
    k = 70
    full = aslam.LargeFilter("ekf", n)
    lo = aslam.LargeFilter("ekf", n, row_begin=0, row_end=k)
    hi = aslam.LargeFilter("ekf", n, row_begin=k, row_end=n)
    for f in (full, lo, hi):
        f.load_state_colmajor(X0, P0)
    for t, (px, py, pth, v, w) in enumerate(_trajectory(4)):
        rng, brg, idx = aslam.large_observations(tx, ty, (px, py, pth), m=8, seed=50 + t)
        a_full = full.step(px, py, np.float32(pth), v, w, 1.0, rng, brg)
        bufs = []
        for f in (lo, hi):
            f.step_phase1(px, py, np.float32(pth), v, w, 1.0, rng, brg)
            f.sync()
            p, c = f.exchange_buffer()
            bufs.append(torch.as_tensor(_DevArray(p, c), device="cuda"))
        total = bufs[0] + bufs[1]
        bufs[0].copy_(total)
        bufs[1].copy_(total)
        torch.cuda.synchronize()
        a_lo, a_hi = lo.step_phase2(), hi.step_phase2()
        assert np.array_equal(a_full, a_lo) and np.array_equal(a_full, a_hi)
        X, P = full.get_state()
        Xl, Pl = lo.get_state()
        Xh, Ph = hi.get_state()
        assert np.array_equal(X, Xl) and np.array_equal(X, Xh)
        assert np.array_equal(P[:k], Pl) and np.array_equal(P[k:], Ph)
    for f in (full, lo, hi):
        f.close()


@pytest.mark.parametrize("cuts", [(70,), (41, 97)])
def test_large_ekf_fused_peer_exchange_one_gpu(aslam, cuts):
    """The fused exchange (peer stores + arrival counters inside the gather / gain kernels, no collective call):
    2 and 3 row shards as handles of ONE process on one GPU, attached through raw block pointers. Each value has
    one owner that stores it into every shard's buffer, so the result is BIT-IDENTICAL to the unsharded handle,
    for several consecutive steps (both buffer parities, growing counters)."""
    tx, ty, X0, G, d = aslam.large_map_state(9, seed=6)
    P0 = aslam.dense_P(G, d)
    n = X0.size
    edges = (0,) + tuple(cuts) + (n,)
    full = aslam.LargeFilter("ekf", n)
    shards = [aslam.LargeFilter("ekf", n, row_begin=edges[i], row_end=edges[i + 1]) for i in range(len(edges) - 1)]
    for f in [full] + shards:
        f.load_state_colmajor(X0, P0)
    blocks = [f.peer_export()[1] for f in shards]
    for r, f in enumerate(shards):
        f.peer_attach(len(shards), r, blocks=blocks)
    for t, (px, py, pth, v, w) in enumerate(_trajectory(5)):
        rng, brg, idx = aslam.large_observations(tx, ty, (px, py, pth), m=8, seed=50 + t)
        a_full = full.step(px, py, np.float32(pth), v, w, 1.0, rng, brg)
        # enqueue every shard before waiting on any of them: a shard's gain kernel spins until all have arrived
        for f in shards:
            f.step_fused(px, py, np.float32(pth), v, w, 1.0, rng, brg, want_assoc=False)
        X, P = full.get_state()
        for i, f in enumerate(shards):
            Xs, Ps = f.get_state()
            assert f.status()[0] == 0
            assert np.array_equal(X, Xs), (t, i)
            assert np.array_equal(P[edges[i]:edges[i + 1]], Ps), (t, i)
    for f in [full] + shards:
        f.close()


def test_large_ekf_fused_peer_exchange_two_gpus(aslam):
    """Same on two physical GPUs of one box (peer access over NVLink, handles of one process): rows [0,k) on
    cuda:0, [k,n) on cuda:1, bit-identical to the unsharded handle. Skipped on a one-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    tx, ty, X0, G, d = aslam.large_map_state(16, seed=6)
    P0 = aslam.dense_P(G, d)
    n = X0.size
    k = n // 2
    full = aslam.LargeFilter("ekf", n, 0)
    shards = [aslam.LargeFilter("ekf", n, 0, row_begin=0, row_end=k), aslam.LargeFilter("ekf", n, 1, row_begin=k, row_end=n)]
    edges = (0, k, n)
    for f in [full] + shards:
        f.load_state_colmajor(X0, P0)
    blocks = [f.peer_export()[1] for f in shards]
    for r, f in enumerate(shards):
        f.peer_attach(2, r, blocks=blocks)
    for t, (px, py, pth, v, w) in enumerate(_trajectory(6)):
        rng, brg, idx = aslam.large_observations(tx, ty, (px, py, pth), m=8, seed=50 + t)
        full.step(px, py, np.float32(pth), v, w, 1.0, rng, brg)
        for f in shards:
            f.step_fused(px, py, np.float32(pth), v, w, 1.0, rng, brg, want_assoc=False)
        X, P = full.get_state()
        for i, f in enumerate(shards):
            Xs, Ps = f.get_state()
            assert f.status()[0] == 0
            assert np.array_equal(X, Xs) and np.array_equal(P[edges[i]:edges[i + 1]], Ps), (t, i)
    for f in [full] + shards:
        f.close()


def test_cfg4_full_size_one_step(aslam, orc):
    """config 4 at BASELINE size: 4096 landmarks, n = 8195, 8 observed (q = 19), two steps against
    the oracle (O(q n^2) on the host cores), plus the symmetry property of the updated P."""
    worst = _run_large(aslam, orc, 64, 8, 2)
    print("cfg4 worst relerr", worst)


def test_large_ekf_fused_step_without_peer_gives_up_cleanly(aslam):
    """A fused step whose peer never arrives: the spin bound expires, the step is ABANDONED (K = 0: predict only,
    the state stays finite), ASLAM_ST_PEER_TIMEOUT is raised and aslam_large_status returns ASLAM_ERR_TIMEOUT.
    And the argument checks come before any launch: a fused step on a handle that was never attached is refused
    and leaves the state untouched."""
    import ctypes as C
    tx, ty, X0, G, d = aslam.large_map_state(9, seed=6)
    P0 = aslam.dense_P(G, d)
    n = X0.size
    k = 70
    lo = aslam.LargeFilter("ekf", n, row_begin=0, row_end=k)
    hi = aslam.LargeFilter("ekf", n, row_begin=k, row_end=n)
    for f in (lo, hi):
        f.load_state_colmajor(X0, P0)
    rng, brg, idx = aslam.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8, seed=50)
    with pytest.raises(aslam.capi.AslamError) as e:  # not attached: refused before the first launch
        lo.step_fused(0.0, 0.0, np.float32(0.0), 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    assert e.value.status == aslam.capi.ASLAM_ERR_UNSUPPORTED
    X, P = lo.get_state()
    assert np.array_equal(X, X0) and np.array_equal(P, P0[:k])
    blocks = [f.peer_export()[1] for f in (lo, hi)]
    for r, f in enumerate((lo, hi)):
        f.peer_attach(2, r, blocks=blocks)
    lo.step_fused(0.0, 0.0, np.float32(0.0), 0.05, 0.1, 1.0, rng, brg, want_assoc=False)  # `hi` never steps
    st = C.c_uint32()
    rc = lo._lib.aslam_large_status(lo._h, C.byref(st), None)
    assert rc == aslam.capi.ASLAM_ERR_TIMEOUT and st.value & (1 << 5)
    X, P = lo.get_state()
    assert np.all(np.isfinite(X)) and np.all(np.isfinite(P))
    assert np.array_equal(X[3:], X0[3:])  # no correction was applied; only the pose moved (predict)
    # P = A P A' + Q: rows / columns 0, 1 and P(2,2) changed, everything else is the loaded matrix
    assert np.array_equal(P[3:, 3:], P0[3:k, 3:])
    for f in (lo, hi):
        f.close()


def _with_env(name, value, fn):
    import os
    old = os.environ.get(name)
    os.environ[name] = value
    try:
        return fn()
    finally:
        if old is None:
            del os.environ[name]
        else:
            os.environ[name] = old


@pytest.mark.parametrize("side,form", [(8, "share"), (13, "share"), (13, "aside")])
def test_large_ekf_overlapped_steps_equal_in_place_steps(aslam, orc, side, form):
    """The overlapped form of the step (solve kernel of step k+1 under the out-of-place update of step k, computing the
    entries of P_k it needs from the stored P_{k-1} and the previous K, H_o P, A) against the in-place, one-stream form,
    on a small map (ASLAM_LARGE_OVERLAP_MIN_BYTES=0 makes the handle overlap whatever its size): X and P BIT-IDENTICAL
    after every one of 7 consecutive messages (both buffer parities, identity start, a get_state between steps that
    does not disturb the pipeline), and the final state within 1e-9 of the oracle's row-subset steps."""
    tx, ty, X0, G, d = aslam.large_map_state(side, seed=6)
    P0 = aslam.dense_P(G, d)
    n = X0.size
    # "share": the solve kernel shares the SMs with the update (large shards); "aside": SMs are set aside for it (short shards)
    ov = _with_env("ASLAM_LARGE_OVERLAP_MIN_BYTES" if form == "share" else "ASLAM_LARGE_RESERVE_MIN_BYTES", "0",
                   lambda: aslam.LargeFilter("ekf", n))
    ip = _with_env("ASLAM_LARGE_NO_OVERLAP", "1", lambda: aslam.LargeFilter("ekf", n))
    for f in (ov, ip):
        f.load_state_colmajor(X0, P0)
    Xo, Po = X0.copy(), P0.copy()
    cfg = orc.default_config()
    a00, a10 = 1.0, 0.0
    for t, (px, py, pth, v, w) in enumerate(_trajectory(7)):
        rng, brg, idx = aslam.large_observations(tx, ty, (px, py, pth), m=8, seed=50 + t)
        a_ov = ov.step(px, py, np.float32(pth), v, w, 1.0, rng, brg)
        a_ip = ip.step(px, py, np.float32(pth), v, w, 1.0, rng, brg)
        q, assoc, a00, a10 = orc.ekf_large_step(cfg, Xo, Po, a00, a10, px, py, np.float32(pth), v, w, 1.0, rng, brg)
        assert np.array_equal(a_ov, a_ip) and np.array_equal(a_ov, assoc)
        if t in (0, 3, 6):
            X1, P1 = ov.get_state()
            X2, P2 = ip.get_state()
            assert np.array_equal(X1, X2) and np.array_equal(P1, P2), t
    X1, P1 = ov.get_state()
    assert ov.status()[0] == 0
    assert relerr(X1, Xo) <= TOL_F64 and relerr(P1, Po.T) <= TOL_F64
    # a load_state in the middle of a pipeline: the loaded matrix is final, the next step starts from it
    ov.load_state_colmajor(X0, P0)
    ip.load_state_colmajor(X0, P0)
    rng, brg, idx = aslam.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8, seed=99)
    for _ in range(2):
        ov.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
        ip.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    X1, P1 = ov.get_state()
    X2, P2 = ip.get_state()
    assert np.array_equal(X1, X2) and np.array_equal(P1, P2)
    ov.close()
    ip.close()
