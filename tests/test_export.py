"""SURVEY 8f #3 / #4, host-side glue (CPU): the headless marker list standing in for visualize_landmark.cpp, the
rosbag-free binary replay, and a syntax check of the ROS1 adapter sources against stub ROS headers (the image has no
ROS: the adapters are shipped as sources, integration/ros1/README.md)."""
import csv
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_marker_list_follows_visualize_landmark(tmp_path, aslam):
    from awesomeslam_b200 import export
    ox, oy = aslam.map_cfg1()
    ex = [2.0000001234, -1.99]
    ey = [0.01, 0.02]
    m = export.marker_list(ox, oy, ex, ey)
    assert [k["id"] for k in m] == list(range(len(ox) + 2))                       # ++id from -1 through both lists
    assert all(k["kind"] == "truth" and (k["g"], k["b"]) == (0.0, 1.0) for k in m[:len(ox)])  # blue ground truth
    assert all(k["kind"] == "estimate" and (k["g"], k["b"]) == (1.0, 0.0) for k in m[len(ox):])  # green estimate
    assert m[len(ox)]["x"] == float(np.float32(ex[0])) and m[len(ox)]["x"] != ex[0]  # const float& in createMarker
    assert all((k["scale_x"], k["scale_y"], k["scale_z"], k["a"], k["frame_id"]) == (0.5, 0.5, 1.0, 1.0, "odom") for k in m)
    export.write_markers_csv(tmp_path / "m.csv", m)
    export.write_markers_json(tmp_path / "m.json", m)
    rows = list(csv.DictReader(open(tmp_path / "m.csv")))
    assert len(rows) == len(m) and rows[-1]["kind"] == "estimate" and float(rows[0]["x"]) == ox[0]
    assert json.load(open(tmp_path / "m.json"))["markers"][3]["id"] == 3


def test_binary_replay_round_trip(tmp_path, aslam):
    from awesomeslam_b200 import export
    lmx, lmy = aslam.map_cfg1()
    rp = aslam.Replay(lmx, lmy, 3, 20, seed0=5)
    p = tmp_path / "run.aslamrp"
    export.save_replay(p, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing)
    r = export.load_replay(p)
    assert (r["batch"], r["steps"], r["obs_per_filter"]) == (3, 20, 8)
    for k, a in (("x", rp.x), ("y", rp.y), ("vx", rp.vx), ("az", rp.az), ("yaw", rp.yaw), ("dt", rp.dt),
                 ("range", rp.range), ("bearing", rp.bearing)):
        assert np.array_equal(r[k].reshape(np.asarray(a).shape), a), k
    assert os.path.getsize(p) == 8 + 16 + 60 * (4 * 8 + 2 * 4) + 60 * 8 * 2 * 4
    with open(p, "r+b") as fh:
        fh.truncate(100)
    try:
        export.load_replay(p)
        assert False, "a truncated file must be rejected"
    except ValueError:
        pass


def test_ros1_adapter_sources_are_valid_cpp():
    """g++ -fsyntax-only against tests/ros_stubs: the adapters use include/aslam_filters.hpp and include/aslam_c.h as
    they are today (no drift between the shipped sources and the headers)."""
    inc = ["-I", os.path.join(ROOT, "tests", "ros_stubs"), "-I", os.path.join(ROOT, "include")]
    for src, extra in (("ekf_node.cpp", []), ("ekf_node.cpp", ["-DASLAM_NODE_UKF"]), ("sensor_landmark_node.cpp", [])):
        subprocess.check_call(["g++", "-std=c++14", "-fsyntax-only"] + extra + inc +
                              [os.path.join(ROOT, "integration", "ros1", src)])
