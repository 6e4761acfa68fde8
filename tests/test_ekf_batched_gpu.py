"""Parity of the batched EKF CUDA path (through the C ABI) against the CPU oracle.

configs 1 and 3 of BASELINE.json. Tolerance: norm-wise relative error <= 1e-9 on X and on P
(FP64 path); N, association indices, wait-list and status must be EQUAL.
"""
import numpy as np
import pytest

from util import TOL_F64, relerr

pytestmark = pytest.mark.gpu


def _run_single(aslam, orc, kind, steps, cfg_over=None, lm=None, check_every=1, vx=0.05, az=0.1, warm=15, seed=1,
                resync_budget=0, blip_tol=1e-5):
    cfg_over = cfg_over or {}
    lmx, lmy = lm if lm is not None else aslam.map_cfg1()
    rp = aslam.Replay(lmx, lmy, 1, steps, warm=warm, vx=vx, az=az, seed0=seed)
    f = aslam.BatchedFilter(kind, 1, 0, **cfg_over)
    o = orc.Filter(kind, **cfg_over)
    worst = {"X": 0.0, "P": 0.0, "resyncs": []}
    for t in range(steps):
        f.set_observations(rp.range[t, 0], rp.bearing[t, 0])
        o.set_observations(rp.range[t, 0], rp.bearing[t, 0])
        f.step(rp.x[t], rp.y[t], rp.yaw[t], rp.vx[t], rp.az[t], rp.dt[t])
        o.step(rp.x[t, 0], rp.y[t, 0], rp.yaw[t, 0], rp.vx[t, 0], rp.az[t, 0], rp.dt[t, 0])
        if t % check_every and t != steps - 1:
            continue
        s = f.get_state()
        X, Z, P = o.state()
        n = o.N
        assert int(s["N"][0]) == n, "state dim differs at step %d" % t
        assert np.array_equal(s["assoc"], o.assoc()), "association differs at step %d" % t
        ex, ep = relerr(s["X"][0, :n], X), relerr(s["P"][0, :n, :n], P)
        if max(ex, ep) > TOL_F64 and len(worst["resyncs"]) < resync_budget and max(ex, ep) <= blip_tol:
            # A float-cast flip: a last-bit difference in sin/cos/atan2 moved a
            # double across a float rounding boundary at one of the reference's FP32 truncation sites,
            # which injects a 6e-8 relative step. Counted, reported, and the CUDA filter is re-seeded
            # from the oracle's state so that every later step is still checked to 1e-9.
            worst["resyncs"].append((t, ex, ep))
            st_before = s["status"].copy()
            f.load_state([n], [X], [Z], [P])
            # the rest of the checkpoint comes from the oracle as well: wait-list with its hit counts, A (EKF)
            wr, wb, wc = o.wait()
            W = aslam.capi.MAX_WAIT
            wait = {"count": np.array([wr.size], np.uint32), "range": np.zeros((1, W), np.float32),
                    "bearing": np.zeros((1, W), np.float32), "hits": np.zeros((1, W), np.uint32)}
            wait["range"][0, :wr.size] = wr
            wait["bearing"][0, :wr.size] = wb
            wait["hits"][0, :wr.size] = wc
            a = o.a_entries() if kind == "ekf" else (None, None)
            f.load_aux(None if a[0] is None else [a[0]], None if a[1] is None else [a[1]], wait=wait, status=st_before)
            continue
        worst["X"] = max(worst["X"], ex)
        worst["P"] = max(worst["P"], ep)
        assert ex <= TOL_F64 and ep <= TOL_F64, "step %d: relerr X %.3e P %.3e" % (t, ex, ep)
        assert not np.any(s["X"][0, n:]) and not np.any(s["P"][0, n:, :]) and not np.any(s["P"][0, :, n:])
        w = f.get_wait()
        wr, wb, wc = o.wait()
        k = int(w["count"][0])
        assert k == wr.size
        assert np.array_equal(w["range"][0, :k], wr) and np.array_equal(w["bearing"][0, :k], wb)
        assert np.array_equal(w["hits"][0, :k], wc)
        if kind == "ekf":
            aux = f.get_aux()
            a00, a10 = o.a_entries()
            assert relerr(aux["Z"][0, :n], Z) <= 1e-15
            assert abs(aux["a00"][0] - a00) <= 1e-12 and abs(aux["a10"][0] - a10) <= 1e-12
        else:
            assert int(f.get_llt_fail()[0]) == o.llt_fail(), "LLT failure column differs at step %d" % t
    f.close()
    return worst, o


def test_cfg1_single_ekf_1000_steps(aslam, orc):
    worst, o = _run_single(aslam, orc, "ekf", 1000)
    assert o.N == 19  # all 8 cylinders of landmarks.yaml mapped
    print("cfg1 worst relerr", worst)


def test_ekf_reference_cap_rejects_batch(aslam, orc):
    """16 landmarks against the reference's cap of 30 (config.h:45): the whole promotion batch is
    dropped (ekf.cpp:263-268) and never retried."""
    worst, o = _run_single(aslam, orc, "ekf", 16, lm=aslam.map_cfg3())
    assert o.N == 3 and o.cap_rejected()


def test_ekf_gated_before_first_observation(aslam):
    f = aslam.BatchedFilter("ekf", 1, 0)
    with pytest.raises(aslam.capi.AslamError) as e:
        f.step([0.0], [0.0], [0.0], [0.0], [0.0], [1.0])
    assert e.value.status == aslam.capi.ASLAM_ERR_NOT_READY
    f.close()


def test_ekf_empty_and_ragged_observations(aslam, orc):
    """Filters of one batch with 0, 3 and 8 observations; empty messages leave the map untouched."""
    lmx, lmy = aslam.map_cfg1()
    steps, B = 40, 3
    rp = aslam.Replay(lmx, lmy, B, steps, seed0=7)
    counts = [0, 3, 8]
    f = aslam.BatchedFilter("ekf", B, 0)
    os_ = [orc.Filter("ekf") for _ in range(B)]
    for t in range(steps):
        rr = [rp.range[t, b, :counts[b]] for b in range(B)]
        bb = [rp.bearing[t, b, :counts[b]] for b in range(B)]
        off = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
        f.set_observations(np.concatenate(rr), np.concatenate(bb), off)
        f.step(rp.x[t], rp.y[t], rp.yaw[t], rp.vx[t], rp.az[t], rp.dt[t])
        for b in range(B):
            os_[b].set_observations(rr[b], bb[b])
            os_[b].step(rp.x[t, b], rp.y[t, b], rp.yaw[t, b], rp.vx[t, b], rp.az[t, b], rp.dt[t, b])
    s = f.get_state()
    for b in range(B):
        X, Z, P = os_[b].state()
        n = os_[b].N
        assert int(s["N"][b]) == n == 3 + 2 * counts[b]
        assert relerr(s["X"][b, :n], X) <= TOL_F64 and relerr(s["P"][b, :n, :n], P) <= TOL_F64
        assert np.array_equal(s["assoc"][off[b]:off[b + 1]], os_[b].assoc())
    f.close()


@pytest.mark.parametrize("msd", [36, 64])
def test_cfg3_batched_ekf_subset(aslam, orc, msd):
    """config 3 at a size the oracle finishes in seconds: 1,024 filters x 16 landmarks x 150 steps,
    per-filter seeds 0x5EED0000 + id, replayed from host buffers in ONE launch (covariance stays
    on-chip between steps). max_state_dim = 36 is the kernel instantiation bench.py times (padded dimension 36, two
    columns per thread, 7 filters per CTA: 147 CTAs here, the last one ragged); 64 the one-column-per-thread one."""
    lmx, lmy = aslam.map_cfg3()
    B, T, L = (1024 if msd == 36 else 256), 150, 16
    rp = aslam.Replay(lmx, lmy, B, T, seed0=0x5EED0000)
    f = aslam.BatchedFilter("ekf", B, 0, max_state_dim=msd)
    assert f.ld == msd
    f.run_host(T, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing, L)
    s = f.get_state()
    cfg = orc.default_config(max_state_dim=msd)
    sec, X, P, N = orc.run_batch("ekf", cfg, B, T, L, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range,
                                 rp.bearing, 0, f.ld)
    assert np.array_equal(s["N"], N) and np.all(N == 35)
    assert not np.any(s["status"])
    ex = [relerr(s["X"][b], X[b]) for b in range(B)]
    ep = [relerr(s["P"][b], P[b]) for b in range(B)]
    print("cfg3 subset worst relerr X %.3e P %.3e" % (max(ex), max(ep)))
    assert max(ex) <= TOL_F64 and max(ep) <= TOL_F64
    f.close()


def test_batched_split_launch_equals_one_launch(aslam):
    """Replaying 40 steps as 1 launch or as 4 x 10 gives bit-identical state (idempotent staging)."""
    lmx, lmy = aslam.map_cfg3()
    B, T, L = 64, 40, 16
    rp = aslam.Replay(lmx, lmy, B, T, seed0=99)
    a = aslam.BatchedFilter("ekf", B, 0, max_state_dim=64)
    b = aslam.BatchedFilter("ekf", B, 0, max_state_dim=64)
    a.run_host(T, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing, L)
    for k in range(0, T, 10):
        sl = slice(k, k + 10)
        b.run_host(10, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl], rp.range[sl],
                   rp.bearing[sl], L)
        b.sync()
    sa, sb = a.get_state(), b.get_state()
    assert np.array_equal(sa["X"], sb["X"]) and np.array_equal(sa["P"], sb["P"]) and np.array_equal(sa["N"], sb["N"])
    a.close(); b.close()


@pytest.mark.parametrize("kind", ["ekf", "ukf"])
def test_handles_on_two_devices_of_one_process(aslam, kind):
    """One process holding handles on cuda:0 and cuda:1 (the shared-memory opt-in of a kernel is per device): the
    same replay gives bit-identical state on both. Skipped on a one-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lmx, lmy = aslam.map_cfg1()  # 8 landmarks: inside the reference's cap of 30 for both kinds
    B, T, L = 48, 20, 8
    rp = aslam.Replay(lmx, lmy, B, T, seed0=21)
    out = []
    for dev in (0, 1):
        f = aslam.BatchedFilter(kind, B, dev, max_state_dim=36 if kind == "ekf" else 30)  # UKF: the reference's cap
        f.run_host(T, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing, L)
        out.append(f.get_state())
        f.close()
    assert np.array_equal(out[0]["X"], out[1]["X"], equal_nan=True)
    assert np.array_equal(out[0]["P"], out[1]["P"], equal_nan=True)
    assert np.array_equal(out[0]["N"], out[1]["N"]) and int(out[0]["N"].max()) == 19


def test_kernel_instantiations_agree_bit_for_bit(aslam):
    """The same replay through the <36,2>, <48,2> and <64,1> instantiations of the batched kernel (18, 24 and 64 threads
    per filter; 7, 5 and 2 filters per CTA): X, P and N identical to the last bit, from reset through promotion and
    growth. The arithmetic per element does not depend on the mapping, so any difference would be an ordering bug."""
    lmx, lmy = aslam.map_cfg3()
    B, T, L = 1024, 48, 16
    rp = aslam.Replay(lmx, lmy, B, T, seed0=0x5EED0000)
    out = {}
    for msd in (36, 48, 64):
        f = aslam.BatchedFilter("ekf", B, 0, max_state_dim=msd)
        assert f.ld == msd
        f.run_host(T, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing, L)
        st = f.get_state()
        out[msd] = (st["X"][:, :35].copy(), st["P"][:, :35, :35].copy(), st["N"].copy())
        f.close()
    assert np.all(out[36][2] == 35)
    for msd in (48, 64):
        for a, b in zip(out[36], out[msd]):
            assert np.array_equal(a, b), msd


def test_pipelined_host_path_equals_one_launch(aslam):
    """run_host without a sync in between (double-buffered staging, copy stream) plus the pipelined
    read_state_async / read_wait: every intermediate X equals the one a synchronous replay of the same
    prefix gives, and the final state is bit-identical to the single-launch replay."""
    lmx, lmy = aslam.map_cfg3()
    B, T, L, C = 96, 48, 16, 6
    rp = aslam.Replay(lmx, lmy, B, T, seed0=7)
    a = aslam.BatchedFilter("ekf", B, 0, max_state_dim=36)
    b = aslam.BatchedFilter("ekf", B, 0, max_state_dim=36)
    ref = aslam.BatchedFilter("ekf", B, 0, max_state_dim=36)
    a.run_host(T, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing, L)
    outs = [np.zeros((B, b.ld)) for _ in range(T // C)]
    tickets = []
    for i, k in enumerate(range(0, T, C)):
        sl = slice(k, k + C)
        b.run_host(C, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl], rp.range[sl], rp.bearing[sl], L)
        tickets.append(b.read_state_async(outs[i]))
    for i, k in enumerate(range(0, T, C)):
        b.read_wait(tickets[i])
        sl = slice(k, k + C)
        ref.run_host(C, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl], rp.range[sl], rp.bearing[sl], L)
        assert np.array_equal(outs[i], ref.get_state(want_P=False, want_assoc=False)["X"]), i
    b.sync()
    sa, sb = a.get_state(), b.get_state()
    assert np.array_equal(sa["X"], sb["X"]) and np.array_equal(sa["P"], sb["P"]) and np.array_equal(sa["N"], sb["N"])
    assert np.array_equal(outs[-1], sb["X"])
    a.close(); b.close(); ref.close()


def test_single_filter_class_mirror(aslam, orc):
    """The reference-named class: cbSensorLandmark / cbOdom, Landmarks out (common.h:93-108)."""
    lmx, lmy = aslam.map_cfg1()
    rp = aslam.Replay(lmx, lmy, 1, 30, seed0=3)
    e = aslam.EKFSlam()
    o = orc.Filter("ekf")
    assert e.cbOdom(aslam.Odometry()) is None  # gated until the first sensor message (ekf.cpp:76)
    for t in range(30):
        e.cbSensorLandmark(rp.range[t, 0].astype(np.float64), rp.bearing[t, 0].astype(np.float64))
        yaw = float(rp.yaw[t, 0])
        msg = aslam.Odometry(rp.x[t, 0], rp.y[t, 0], np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2), rp.vx[t, 0], rp.az[t, 0])
        out = e.cbOdom(msg, 1.0)
        o.set_observations(rp.range[t, 0], rp.bearing[t, 0])
        yaw32 = aslam.capi.lib().aslam_quat2euler(msg.qw, 0.0, 0.0, msg.qz)
        o.step(msg.x, msg.y, yaw32, msg.vx, msg.az, 1.0)
    X, Z, P = o.state()
    lx, ly = out
    assert lx.size == (o.N - 3) // 2 == 8
    assert relerr(lx, X[3::2]) <= TOL_F64 and relerr(ly, X[4::2]) <= TOL_F64
    e.close()


@pytest.mark.parametrize("kind", ["ekf", "ukf"])
def test_checkpoint_restore_continues_bit_identically(aslam, kind):
    """Checkpoint in the middle of the promotion phase (wait-list half counted, for the EKF a non-trivial A kept over a
    zero-twist message), restore into a FRESH handle with load_state + load_aux, continue: bit-identical to the
    uninterrupted run. And load_state alone on a USED handle resets the wait-list, A and the status."""
    lmx, lmy = aslam.map_cfg1()
    B, T, L, cut = 8, 40, 8, 6
    rp = aslam.Replay(lmx, lmy, B, T, warm=4, vx=0.02, seed0=991)   # moving from message 4 on: A != I at the cut
    rp.vx[cut + 1] = 0.0                                   # a message that keeps the previous A (ekf.cpp:206)

    def run(f, t0, t1):
        for t in range(t0, t1):
            f.set_observations([rp.range[t, b] for b in range(B)], [rp.bearing[t, b] for b in range(B)])
            f.step(rp.x[t], rp.y[t], rp.yaw[t], rp.vx[t], rp.az[t], rp.dt[t])

    a = aslam.BatchedFilter(kind, B, 0)
    run(a, 0, cut)
    s, aux, w = a.get_state(), a.get_aux(), a.get_wait()
    assert int(w["count"].max()) > 0 and int(w["hits"].max()) < 10  # nothing promoted yet: the counts matter
    b = aslam.BatchedFilter(kind, B, 0)
    b.load_state(s["N"], s["X"], aux["Z"], s["P"])
    b.load_aux(aux["a00"] if kind == "ekf" else None, aux["a10"] if kind == "ekf" else None, wait=w,
               status=s["status"])
    run(a, cut, T)
    run(b, cut, T)
    sa, sb = a.get_state(), b.get_state()
    assert np.array_equal(sa["N"], sb["N"]) and int(sa["N"].min()) > 3
    assert np.array_equal(sa["X"], sb["X"]) and np.array_equal(sa["P"], sb["P"])
    wa, wb = a.get_wait(), b.get_wait()
    assert all(np.array_equal(wa[k], wb[k]) for k in wa)
    # a used handle: load_state puts the wait-list, A and the status back to their initial values
    a.load_state(s["N"], s["X"], aux["Z"], s["P"])
    w0, x0 = a.get_wait(), a.get_aux()
    assert not w0["count"].any() and not a.get_state(want_P=False, want_assoc=False)["status"].any()
    if kind == "ekf":
        assert np.all(x0["a00"] == 1.0) and np.all(x0["a10"] == 0.0)
    a.close()
    b.close()


def test_set_observations_rejects_decreasing_offsets(aslam):
    f = aslam.BatchedFilter("ekf", 2, 0)
    import ctypes as C
    off = np.array([0, 3, 2], np.uint32)
    r = np.zeros(4, np.float32)
    rc = f._lib.aslam_filter_set_observations(f._h, aslam.capi.ptr(r), aslam.capi.ptr(r), aslam.capi.ptr(off))
    assert rc == aslam.capi.ASLAM_ERR_ARG
    f.close()


def test_cfg3_long_free_run_blip_count(aslam, orc):
    """SURVEY 7.3 item 1 / VERDICT r1 (parity 3): how many of N free-running cfg3 filters leave the 1e-9 band over 1000
    messages, and by how much (a last-bit sin/cos/atan2 difference between CUDA and glibc can flip one of the
    reference's float casts). 1,024 filters x 1000 messages, never re-seeded; the CPU side is the compiled reference
    (oracle/_ref, cap lifted to 64) when built, else the restated oracle. Count and maximum are printed; the bar: no
    filter beyond 1e-9."""
    from oracle import ref as R
    lmx, lmy = aslam.map_cfg3()
    B, T, L = 1024, 1000, 16
    rp = aslam.Replay(lmx, lmy, B, T, seed0=0x5EED0000)
    f = aslam.BatchedFilter("ekf", B, 0, max_state_dim=36)
    for t0 in range(0, T, 125):
        sl = slice(t0, t0 + 125)
        f.run_host(125, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl], rp.range[sl], rp.bearing[sl], L)
    s = f.get_state()
    if R.available("ekf", True):
        sec, X, P, N = R.run_batch("ekf", B, T, L, rp.x, rp.y, rp.qw, rp.qz, rp.vx, rp.az, rp.range, rp.bearing, 0, f.ld,
                                   cap64=True)
    else:
        sec, X, P, N = orc.run_batch("ekf", orc.default_config(max_state_dim=64), B, T, L, rp.x, rp.y, rp.yaw, rp.vx,
                                     rp.az, rp.dt, rp.range, rp.bearing, 0, f.ld)
    assert np.array_equal(s["N"], N) and not np.any(s["status"])
    err = np.array([max(relerr(s["X"][b], X[b]), relerr(s["P"][b], P[b])) for b in range(B)])
    beyond = int(np.sum(err > TOL_F64))
    print("cfg3 free run, %d filters x %d messages: %d beyond 1e-9, worst %.3e" % (B, T, beyond, err.max()))
    assert beyond == 0
    f.close()


def _far_apart_observations(count, skip=0):
    """(range, bearing) float32 pairs of `count` points on a 1 m grid around a robot at rest at the origin (so every
    pair of points is at least 1 m apart, twice the association threshold), none closer than 2 m to the robot."""
    pts = [(i, j) for i in range(-12, 13) for j in range(-12, 13) if i * i + j * j >= 4]
    pts = np.array(pts[skip:skip + count], dtype=np.float64)
    assert pts.shape[0] == count
    return np.hypot(pts[:, 0], pts[:, 1]).astype(np.float32), np.arctan2(pts[:, 1], pts[:, 0]).astype(np.float32)


@pytest.mark.parametrize("kind", ["ekf", "ukf"])
def test_wait_list_and_observation_bounds(aslam, orc, kind):
    """ASLAM_MAX_OBS = 64 observations per message and ASLAM_MAX_WAIT = 256 wait-list entries bound what the reference
    leaves unbounded (std::vector, ekf.h:104). AT the bound the kernels equal the oracle: 4 messages x 64 new, mutually
    distant observations fill the wait-list to exactly 256 entries with a clean status. BEYOND it they say so: one more
    new observation raises ASLAM_ST_WAIT_OVERFLOW (the list keeps its 256 entries), a message of 65 observations raises
    ASLAM_ST_OBS_OVERFLOW (the first 64 are used)."""
    f = aslam.BatchedFilter(kind, 1, 0)
    o = orc.Filter(kind)
    odom = ([0.0], [0.0], [0.0], [0.0], [0.0], [1.0])
    for msg in range(4):
        r, b = _far_apart_observations(64, skip=64 * msg)
        f.set_observations(r, b)
        f.step(*odom)
        o.set_observations(r, b)
        o.step(0.0, 0.0, 0.0, 0.0, 0.0, 1.0)
    w = f.get_wait()
    wr, wb, wh = o.wait(cap=512)
    s = f.get_state()
    assert int(w["count"][0]) == 256 == wr.size and int(s["status"][0]) == 0 and int(s["N"][0]) == 3 == o.N
    assert np.array_equal(w["range"][0], wr) and np.array_equal(w["bearing"][0], wb) and np.array_equal(w["hits"][0], wh)
    assert np.all(wh == 1)
    # a re-observation of an entry is a hit, not an append: still no overflow
    r, b = _far_apart_observations(1, skip=100)
    f.set_observations(r, b)
    f.step(*odom)
    w = f.get_wait()
    assert int(f.get_state()["status"][0]) == 0 and int(w["count"][0]) == 256 and int(w["hits"][0, 100]) == 2
    # entry 257
    r, b = _far_apart_observations(1, skip=256)
    f.set_observations(r, b)
    f.step(*odom)
    st = int(f.get_state()["status"][0])
    assert st & aslam.capi.ST_WAIT_OVERFLOW and not st & aslam.capi.ST_OBS_OVERFLOW
    assert int(f.get_wait()["count"][0]) == 256
    f.close()
    # 65 observations in one message
    g = aslam.BatchedFilter(kind, 1, 0)
    r, b = _far_apart_observations(65)
    g.set_observations(r, b)
    g.step(*odom)
    st = int(g.get_state()["status"][0])
    assert st & aslam.capi.ST_OBS_OVERFLOW
    w = g.get_wait()
    assert int(w["count"][0]) == 64 and np.array_equal(w["range"][0, :64], r[:64])
    g.close()
