"""Laser front end: sensor_landmark.cpp restated in oracle/laser_front.hpp, cross-checked against
the independent numpy restatement (CPU), and the CUDA kernel behind aslam_laser_* against the oracle (GPU).
Tolerance: landmark counts equal; range relative / bearing absolute difference <= 2e-6 (the published values are
floats; the two sides run different eigen-solvers / libm, so a last-bit float difference is allowed)."""
import numpy as np
import pytest

LASER_TOL = 2e-6


def _wall():
    """41 rays (100..140 degrees) on a straight wall 1.5 m from the sensor."""
    return (1.5 / np.cos(np.deg2rad(np.arange(100, 141) - 120.0))).astype(np.float32)


def _scans(aslam, B, seed, cfg3=False):
    lmx, lmy = aslam.map_cfg3() if cfg3 else aslam.map_cfg1()
    poses = aslam.laser_poses(B, seed=seed, **(aslam.replay.LASER_POSES_CFG3 if cfg3 else {}))
    return aslam.laser_scans(lmx, lmy, poses, seed=seed + 1), poses, lmx, lmy


def test_generator_and_oracle_find_the_cylinders(aslam, orc):
    """Synthetic LDS scans of the landmarks.yaml map: every cylinder within the 3.5 m range is found, and its fitted
    centre is within 3 cm of the truth (range noise sigma 1 cm)."""
    sc, poses, lmx, lmy = _scans(aslam, 32, 11)
    rg, bg, cnt, _ = orc.laser_extract(sc)
    assert np.all(cnt >= 3)
    for i in range(32):
        d = np.hypot(np.array(lmx) - poses[i, 0], np.array(lmy) - poses[i, 1])
        b = np.arctan2(np.array(lmy) - poses[i, 1], np.array(lmx) - poses[i, 0]) - poses[i, 2]
        for k in range(cnt[i]):
            ex = poses[i, 0] + rg[i, k] * np.cos(poses[i, 2] + bg[i, k])
            ey = poses[i, 1] + rg[i, k] * np.sin(poses[i, 2] + bg[i, k])
            assert np.min(np.hypot(np.array(lmx) - ex, np.array(lmy) - ey)) < 0.10
        assert d.size == b.size


def test_oracle_matches_numpy_restatement(aslam, orc):
    from oracle import np_restatement as R
    sc, _, _, _ = _scans(aslam, 24, 3)
    sc[5, 100:141] = _wall()  # a straight wall: a long cluster the circle test must reject
    sc[6, :] = np.inf     # empty scan
    rg, bg, cnt, _ = orc.laser_extract(sc)
    assert cnt[6] == 0
    for i in range(sc.shape[0]):
        ref = R.laser_extract(sc[i])
        assert ref is not None and len(ref) == cnt[i], i
        for k, (r, b) in enumerate(ref):
            assert abs(r - rg[i, k]) <= LASER_TOL * abs(r) and abs(b - bg[i, k]) <= LASER_TOL


def test_first_and_last_ray_on_one_object_is_the_reference_assert(aslam, orc):
    """sensor_landmark.cpp:63-82: when ray 0 and ray 359 fall on the same object the first-cluster loop swallows the
    scan and assert(theta < 359) fires; the oracle reports -1."""
    from oracle import np_restatement as R
    sc = np.full((1, 360), np.inf, np.float32)
    sc[0, :6] = 2.0
    sc[0, 355:] = 2.0
    assert orc.laser_extract(sc)[2][0] == -1
    assert R.laser_extract(sc[0]) is None


@pytest.mark.gpu
@pytest.mark.parametrize("cfg3", [False, True])
def test_laser_kernel_matches_oracle(aslam, orc, cfg3):
    sc, _, _, _ = _scans(aslam, 512, 21, cfg3=cfg3)
    sc[7, :] = np.inf
    sc[8, :6] = 2.0
    sc[8, 355:] = 2.0
    sc[9, 100:141] = _wall()
    lm = aslam.LandMarks(max_scans=512)
    rg, bg, cnt = lm.callback(sc)
    org, obg, ocnt, _ = orc.laser_extract(sc)
    assert np.array_equal(cnt, ocnt)
    assert cnt[7] == 0 and cnt[8] == -1
    worst_r = worst_b = 0.0
    for i in range(sc.shape[0]):
        for k in range(max(cnt[i], 0)):
            worst_r = max(worst_r, abs(rg[i, k] - org[i, k]) / abs(org[i, k]))
            worst_b = max(worst_b, abs(bg[i, k] - obg[i, k]))
    print("laser kernel vs oracle: %d landmarks, worst range rel %.2e, bearing abs %.2e" % (int(cnt[cnt > 0].sum()), worst_r, worst_b))
    assert worst_r <= LASER_TOL and worst_b <= LASER_TOL
    lm.close()


@pytest.mark.gpu
def test_laser_to_filter_pipeline(aslam, orc):
    """Scan in, map out: the kernel's (range, bearing) lists drive the single-filter EKF mirror exactly as the
    oracle's lists drive the oracle filter (cbSensorLandmark then cbOdom, robot at rest for 14 messages)."""
    lmx, lmy = aslam.map_cfg1()
    pose = np.array([[0.05, -0.02, np.deg2rad(33.0)]])
    th = float(pose[0, 2])
    od = aslam.Odometry(x=pose[0, 0], y=pose[0, 1], qw=np.cos(th / 2), qz=np.sin(th / 2))
    yaw = aslam.capi.lib().aslam_quat2euler(od.qw, od.qx, od.qy, od.qz)
    lm = aslam.LandMarks(max_scans=1)
    e = aslam.EKFSlam()
    o = orc.Filter("ekf")
    n = 0
    for t in range(14):
        sc = aslam.laser_scans(lmx, lmy, pose, seed=100 + t)
        rg, bg, cnt = lm.callback(sc)
        org, obg, ocnt, _ = orc.laser_extract(sc)
        n = int(cnt[0])
        assert n == int(ocnt[0]) and n >= 3
        e.cbSensorLandmark(rg[0, :n].astype(np.float64), bg[0, :n].astype(np.float64))
        e.cbOdom(od)
        o.set_observations(org[0, :n], obg[0, :n])
        o.step(pose[0, 0], pose[0, 1], yaw, 0.0, 0.0, np.float32(1.0))
    st = e.state()
    Xo, Zo, Po = o.state()
    assert st["N"] == o.N == 3 + 2 * n and st["status"] == 0
    assert np.max(np.abs(st["X"] - Xo[:st["N"]])) <= 1e-5  # float-level differences of the observations, 14 updates
    # the mapped cylinders sit where the world has them (robot pose known from odometry), up to the +1 degree rotation
    # the reference's front end applies to every bearing (sensor_landmark.cpp:87,98 as GCC evaluates it: 0.035 m at 2 m)
    for l in range(n):
        d = np.hypot(np.array(lmx) - st["X"][3 + 2 * l], np.array(lmy) - st["X"][4 + 2 * l])
        assert d.min() < 0.08
    lm.close()
    e.close()


@pytest.mark.gpu
def test_synthetic_scan_kernel_matches_numpy_generator(aslam):
    """the CUDA scan generator (one thread per ray, ray-cylinder intersection) against the numpy
    generator on the same poses and the same noise array: identical hit pattern, ranges equal to one float ulp."""
    lmx, lmy = aslam.map_cfg3()
    B = 256
    poses = aslam.laser_poses(B, seed=9, **aslam.replay.LASER_POSES_CFG3)
    noise = np.random.default_rng(10).standard_normal((B, 360)).astype(np.float32)
    ref = aslam.laser_scans(lmx, lmy, poses, noise=noise)
    lm = aslam.LandMarks(max_scans=B)
    got = lm.synthesize(lmx, lmy, poses, noise=noise)
    assert np.array_equal(np.isfinite(got), np.isfinite(ref)) and np.isfinite(ref).sum() > 50 * B
    f = np.isfinite(ref)
    assert np.max(np.abs(got[f] - ref[f]) / ref[f]) <= 1.2e-7
    # and the scans it makes are good scans: the front end finds the cylinders in range
    rg, bg, cnt = lm.callback(got)
    assert np.all(cnt >= 3)
    lm.close()


@pytest.mark.gpu
def test_laser_device_path_batches_larger_than_the_work_list(aslam):
    """aslam_laser_extract_device with more scans than the handle was created for (its dense work list of accepted
    clusters holds max_scans scans: larger batches are processed in chunks) equals the host-buffer path of a handle
    that is large enough."""
    import torch
    sc, _, _, _ = _scans(aslam, 200, 5, cfg3=True)
    big = aslam.LandMarks(max_scans=200)
    rg, bg, cnt = big.callback(sc)
    big.close()
    small = aslam.LandMarks(max_scans=48)
    dev = torch.device("cuda", 0)
    d_sc = torch.from_numpy(sc).to(dev)
    d_rg = torch.zeros((200, 32), dtype=torch.float32, device=dev)
    d_bg = torch.zeros((200, 32), dtype=torch.float32, device=dev)
    d_cnt = torch.zeros(200, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    small.callback_device(d_sc.data_ptr(), 200, d_rg.data_ptr(), d_bg.data_ptr(), d_cnt.data_ptr())
    small.sync()
    assert small.launch_count == 2 * 5  # ceil(200 / 48) chunks, classification + fit each
    assert np.array_equal(d_cnt.cpu().numpy(), cnt)
    for i in range(200):
        k = max(int(cnt[i]), 0)
        assert np.array_equal(d_rg[i, :k].cpu().numpy(), rg[i, :k]) and np.array_equal(d_bg[i, :k].cpu().numpy(), bg[i, :k])
    small.close()
