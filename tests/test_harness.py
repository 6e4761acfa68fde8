"""The ROS-free C++ replay harness (awesomeslam_b200/harness/replay_main.cpp over include/aslam_filters.hpp): the
reference-style callback classes aslam::EKFSlam / aslam::UKFSlam driven message by message, and the batched entry."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _harness():
    from awesomeslam_b200 import build
    build.build_library()
    return build.build_harness()


def _run(args):
    return subprocess.run([_harness()] + args, cwd=ROOT, capture_output=True, text=True, timeout=300)


def test_harness_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run(["--steps", "5"])
    assert r.returncode == 1 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_harness_single_ekf_maps_the_yaml_cylinders(tmp_path):
    csv = str(tmp_path / "track.csv")
    r = _run(["--filter", "ekf", "--steps", "300", "--csv", csv])
    assert r.returncode == 0, r.stderr
    m = re.search(r"mapped (\d+) of (\d+) landmarks, worst error ([0-9.]+) m", r.stdout)
    assert m and m.group(1) == "8" and m.group(2) == "8" and float(m.group(3)) < 0.02, r.stdout
    assert "state dim N = 19" in r.stdout
    rows = open(csv).read().strip().split("\n")
    assert len(rows) == 300 and len(rows[-1].split(",")) == 1 + 16


@pytest.mark.gpu
def test_harness_batched_and_ukf():
    r = _run(["--filter", "ekf", "--steps", "100", "--batch", "512", "--seed", "99"])
    assert r.returncode == 0 and "state dim N = 19" in r.stdout and "status 0x0" in r.stdout, r.stdout + r.stderr
    r = _run(["--filter", "ukf", "--steps", "40"])
    assert r.returncode == 0 and "ukf-slam: 1 filter(s) x 40 messages" in r.stdout, r.stdout + r.stderr
