"""The pin: the reference's OWN sources (ekf.cpp, ukf.cpp, sensor_landmark.cpp and the awesome_slam headers, compiled
unmodified into oracle/_ref by oracle/ref_shim/Makefile) against the restated oracle that every GPU parity test uses.

Everything here must be BIT-EQUAL: the two share only the dense-matrix arithmetic of oracle/dense.hpp (Eigen itself is
absent from the image and its kernels are not bit-reproducible, see oracle/ref_shim/include/aslam_ref_shim/
eigen_dense.hpp); control flow, cast sites, association, wait-list, growth, the order of every matrix expression and
the ROS callback shell are the reference's text on one side and the restatement on the other.

CPU only. oracle/_ref is built here from /root/reference; on the GPU box the prebuilt libraries travel with the
snapshot (nothing reads /root/reference at run time).
"""
import os

import numpy as np
import pytest

from oracle import ref as R

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref neither prebuilt nor buildable here")


@pytest.fixture(scope="module")
def ref():
    R.build()
    return R


def _dt_like_reference(now, last_time_f32):
    """ekf.cpp:80: float delta_time = std::min(now - last_time, 1.0) with a float last_time member."""
    return np.float32(min(float(now) - float(last_time_f32), 1.0))


def _drive(kind, orc, ref, msgs, cap64=False):
    """Feeds the same message sequence to a reference node and to the oracle; compares after every message.
    msgs: iterable of dicts {t, obs=(range64, bearing64) or None, odom=(x, y, qw, qx, qy, qz, vx, az) or None}."""
    node = ref.Node(kind, cap64=cap64)
    o = orc.Filter(kind, max_state_dim=64 if cap64 else 30)
    last_time = np.float32(0.0)  # initialize(): last_time = ros::Time::now() at construction (clock 0)
    n_odom = 0
    for k, m in enumerate(msgs):
        node.set_time(m["t"])
        if m.get("obs") is not None:
            rng64, brg64 = m["obs"]
            node.landmarks(rng64, brg64)
            # ekf.cpp:111 data.assign(msg->x[i], msg->y[i]): the message doubles are rounded to float
            o.set_observations(np.asarray(rng64, np.float64).astype(np.float32),
                               np.asarray(brg64, np.float64).astype(np.float32))
        if m.get("odom") is not None:
            x, y, qw, qx, qy, qz, vx, az = m["odom"]
            init_z = node.flags[1]
            node.odom(x, y, qw, qx, qy, qz, vx, az)
            if not init_z:  # the reference returns before touching last_time while init_z is set (ekf.cpp:76-77)
                dt = _dt_like_reference(m["t"], last_time)
                last_time = np.float32(m["t"])
                yaw = ref.quat2euler(qw, qx, qy, qz, kind)  # tools.h:62-66, the reference's own function
                stepped = o.step(x, y, yaw, vx, az, dt)
                assert stepped
                n_odom += 1
            else:
                assert not o.step(x, y, 0.0, vx, az, 1.0)
        Xr, Zr, Pr = node.state()
        Xo, Zo, Po = o.state()
        assert node.N == o.N, (k, node.N, o.N)
        assert np.array_equal(Xr, Xo), (k, np.abs(Xr - Xo).max())
        assert np.array_equal(Zr, Zo), (k, np.abs(Zr - Zo).max())
        assert np.array_equal(Pr, Po), (k, np.abs(Pr - Po).max())
        wr, wo = node.wait(), o.wait(cap=1024)
        for a, b in zip(wr, wo):
            assert np.array_equal(a, b), (k, a, b)
        if kind == "ekf":
            assert node.a_entries() == o.a_entries(), k
        px, py, total = node.published()
        assert total == n_odom, (k, total, n_odom)
        if n_odom:  # convertToLandmarkMsg (common.h:93-108) of the state just compared
            assert np.array_equal(px, Xo[3::2]) and np.array_equal(py, Xo[4::2]), k
    return node, o


def _replay_msgs(rp, f=0):
    for t in range(rp.steps):
        yield dict(t=float(t + 1), obs=(rp.range[t, f].astype(np.float64), rp.bearing[t, f].astype(np.float64)),
                   odom=(rp.x[t, f], rp.y[t, f], rp.qw[t, f], 0.0, 0.0, rp.qz[t, f], rp.vx[t, f], rp.az[t, f]))


def test_ref_libraries_are_the_reference_as_shipped(ref):
    assert ref.lib("ekf").ref_max_state_dim() == 30 and ref.lib("ukf").ref_max_state_dim() == 30  # config.h:45
    assert ref.lib("ekf", cap64=True).ref_max_state_dim() == 64  # the one build-time constant change (Makefile)
    # the recipe compiles the reference from where it lies; no reference source text is kept in the repository
    mk = open(os.path.join(os.path.dirname(R.__file__), "ref_shim", "Makefile")).read()
    assert "/root/reference/awesome_slam" in mk and "$(REF)/src/ekf/ekf.cpp" in mk


def test_cfg1_ekf_1000_messages_bit_equal(aslam_host, orc, ref):
    lmx, lmy = aslam_host.map_cfg1()
    rp = aslam_host.Replay(lmx, lmy, 1, 1000, seed0=aslam_host.replay.SEED_CFG1)
    node, o = _drive("ekf", orc, ref, _replay_msgs(rp))
    assert node.N == 19
    # the yaw the replay hands to the CUDA path is the reference's quat2euler of the same quaternion
    for t in (0, 17, 400, 999):
        assert np.float32(ref.quat2euler(rp.qw[t, 0], 0.0, 0.0, rp.qz[t, 0])) == rp.yaw[t, 0]


def test_cfg2_ukf_1000_messages_bit_equal(aslam_host, orc, ref):
    """Includes every LLT failure of the free-running reference UKF (Eigen's half-factored L, restated in
    oracle/dense.hpp and used by both sides) and its cap rejections (N reaches 29)."""
    lmx, lmy = aslam_host.map_cfg1()
    rp = aslam_host.Replay(lmx, lmy, 1, 1000, seed0=aslam_host.replay.SEED_CFG1)
    node, o = _drive("ukf", orc, ref, _replay_msgs(rp))
    assert node.N == o.N


@pytest.mark.parametrize("kind", ["ekf", "ukf"])
def test_irregular_message_stream_bit_equal(aslam_host, orc, ref, kind):
    """What the replay never does: odometry before the first landmark message (init_z gate), empty and ragged
    observation lists, several odometry messages per landmark message (the list persists, ekf.cpp:105), clock steps
    below and above one second (delta_time clamp, ekf.cpp:80), zero twist (A keeps its previous value,
    ekf.cpp:206), bearings outside (-pi, pi] (normalizeAngle in place, ekf.cpp:150), a full quaternion."""
    lmx, lmy = aslam_host.map_cfg1()
    rp = aslam_host.Replay(lmx, lmy, 1, 160, seed0=77)
    g = np.random.default_rng(5)
    msgs, t = [], 0.0
    msgs.append(dict(t=0.25, odom=(0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.1, 0.1)))  # gated: no landmark message yet
    for k in range(160):
        t += float(g.choice([0.25, 0.5, 1.0, 1.75]))
        r = rp.range[k, 0].astype(np.float64)
        b = rp.bearing[k, 0].astype(np.float64)
        keep = g.random(8) < (1.0 if k < 20 else 0.7)
        if k == 30:
            keep[:] = False  # empty list
        b = b + 2 * np.pi * g.integers(-1, 2, 8)  # same direction, unnormalised
        m = dict(t=t)
        if k % 3 != 2:
            m["obs"] = (r[keep], b[keep])
        tilt = 0.02 * g.standard_normal(2)  # small roll/pitch: x, y components of the quaternion
        q = np.array([rp.qw[k, 0], tilt[0], tilt[1], rp.qz[k, 0]])
        q /= np.linalg.norm(q)
        vx, az = rp.vx[k, 0], rp.az[k, 0]
        if k % 11 == 5:
            az = 0.0
        if k % 13 == 7:
            vx = 0.0
        if k % 17 == 3:
            az = 0.0005  # below the 1e-3 straight-line switch of common.h:52
        m["odom"] = (rp.x[k, 0], rp.y[k, 0], q[0], q[1], q[2], q[3], vx, az)
        msgs.append(m)
    node, o = _drive(kind, orc, ref, msgs)
    assert node.N > 3


@pytest.mark.parametrize("kind", ["ekf", "ukf"])
def test_cap_rejection_bit_equal(aslam_host, orc, ref, kind):
    """16 landmarks promoted at once: N = 35 >= MAX_LANDMARK_COUNT, the whole batch is dropped (ekf.cpp:263-268); the
    wait-list counters have moved past MIN_LANDMARK_OCC, so those candidates are never retried."""
    lmx, lmy = aslam_host.map_cfg3()
    rp = aslam_host.Replay(lmx, lmy, 1, 30, seed0=3)
    node, o = _drive(kind, orc, ref, _replay_msgs(rp))
    assert node.N == 3 and o.cap_rejected()
    assert np.all(node.wait()[2][:16] > 10)


@pytest.mark.parametrize("kind", ["ekf", "ukf"])
def test_cfg3_shape_with_cap_lifted_bit_equal(aslam_host, orc, ref, kind):
    """The cfg3 shape (16 landmarks, n = 35) needs MAX_LANDMARK_COUNT lifted -- the only constant changed, at build
    time, in libref_*_cap64.so. Batch runner on both sides (this is also the CPU-baseline code path of bench.py)."""
    lmx, lmy = aslam_host.map_cfg3()
    B, T, L = 6, 60 if kind == "ukf" else 200, 16
    rp = aslam_host.Replay(lmx, lmy, B, T, seed0=aslam_host.replay.SEED_CFG3)
    _, Xr, Pr, Nr = ref.run_batch(kind, B, T, L, rp.x, rp.y, rp.qw, rp.qz, rp.vx, rp.az, rp.range, rp.bearing, 2, 36,
                                  cap64=True)
    cfg = orc.default_config(max_state_dim=64)
    _, Xo, Po, No = orc.run_batch(kind, cfg, B, T, L, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing,
                                  2, 36)
    assert np.array_equal(Nr, No) and np.all(Nr == 35)
    assert np.array_equal(Xr, Xo) and np.array_equal(Pr, Po)


def test_laser_front_end_bit_equal(aslam_host, orc, ref):
    """LandMarks::callback of the compiled reference against oracle/laser_front.hpp on synthetic scans of both maps:
    equal landmark counts, every published float equal. (This comparison is what exposed the argument-evaluation
    order of sensor_landmark.cpp:87,98 -- see oracle/laser_front.hpp.)"""
    from awesomeslam_b200 import replay as RP
    total = 0
    for (mx, my), kw, seed in ((aslam_host.map_cfg1(), {}, 5), (aslam_host.map_cfg3(), RP.LASER_POSES_CFG3, 7)):
        sc = aslam_host.laser_scans(mx, my, aslam_host.laser_poses(1500, seed=seed, **kw), seed=seed + 1)
        sc[3, :] = np.inf          # nothing in range
        sc[4, 350:] = 1.0          # first and last ray on one object: the reference's assert -> -1
        sc[4, :10] = 1.0
        r1, b1, c1, _ = orc.laser_extract(sc)
        r2, b2, c2, _ = ref.laser_extract(sc)
        assert np.array_equal(c1, c2) and c1[3] == 0 and c1[4] == -1
        m = np.arange(r1.shape[1])[None, :] < np.maximum(c1, 0)[:, None]
        assert np.array_equal(r1[m], r2[m]) and np.array_equal(b1[m], b2[m])
        total += int(m.sum())
    assert total > 15000


def test_golden_vectors_are_reference_outputs(ref):
    """tests/golden/cfg1_ekf.npz and cfg2_ukf.npz hold X, P, N at selected messages; they are what the compiled
    reference produces from the stored inputs (make_golden.py asserts the same when it writes them)."""
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    for kind, name in (("ekf", "cfg1_ekf.npz"), ("ukf", "cfg2_ukf.npz")):
        g = np.load(os.path.join(here, name))
        node = ref.Node(kind)
        steps = list(g["steps"])
        for t in range(int(steps[-1]) + 1):
            node.set_time(float(t + 1))
            node.landmarks(g["range"][t].astype(np.float64), g["bearing"][t].astype(np.float64))
            node.odom(g["x"][t], g["y"][t], g["qw"][t], 0.0, 0.0, g["qz"][t], g["vx"][t], g["az"][t])
            if t in steps:
                k = steps.index(t)
                X, Z, P = node.state()
                n = node.N
                assert n == int(g["N"][k])
                assert np.array_equal(X, g["X"][k][:n]) and np.array_equal(P, g["P"][k][:n, :n]), (kind, t)
