"""Maps and the deterministic synthetic replay of SURVEY.md section 8d.

The filter itself never reads landmarks.yaml (in the reference only gazebo_spawn.cpp:46-47 and
visualize_landmark.cpp:40-41 do); it is the input of the replay generator that stands in for
Gazebo's /odom and the sensor_landmark node's (range, bearing) lists.
"""
import os

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LANDMARKS_YAML = os.path.join(_HERE, "config", "landmarks.yaml")


def load_landmarks_yaml(path=LANDMARKS_YAML):
    """Parses the two ROS-param lists `landmarks/x`, `landmarks/y` (flow or block style)."""
    import yaml
    with open(path) as f:
        d = yaml.safe_load(f)
    x = np.asarray(d["landmarks/x"], np.float64)
    y = np.asarray(d["landmarks/y"], np.float64)
    assert x.shape == y.shape
    return x, y


def map_cfg1():
    """configs 1-2: the 8 cylinders of landmarks.yaml."""
    return load_landmarks_yaml()


def map_cfg3():
    """config 3: 16 landmarks = the yaml map + 8 more, pairwise >= 1.58 m apart."""
    x, y = load_landmarks_yaml()
    ex = np.array([4, -4, 0, 0, 1.5, -1.5, 1.5, -1.5], np.float64)
    ey = np.array([0, 0, 4, -4, 1.5, 1.5, -1.5, -1.5], np.float64)
    return np.concatenate([x, ex]), np.concatenate([y, ey])


class Replay:
    """Step-major message arrays: odometry [steps][batch], observations [steps][batch][L]."""

    def __init__(self, lm_x, lm_y, batch, steps, warm=15, vx=0.05, az=0.1, seed0=1, first_filter=0):
        lm_x = capi.as_np(lm_x, np.float64)
        lm_y = capi.as_np(lm_y, np.float64)
        L = lm_x.size
        self.L, self.batch, self.steps = L, batch, steps
        self.lm_x, self.lm_y = lm_x, lm_y
        self.x = np.zeros((steps, batch), np.float64)
        self.y = np.zeros((steps, batch), np.float64)
        self.yaw = np.zeros((steps, batch), np.float32)
        self.vx = np.zeros((steps, batch), np.float64)
        self.az = np.zeros((steps, batch), np.float64)
        self.dt = np.zeros((steps, batch), np.float32)
        self.range = np.zeros((steps, batch, L), np.float32)
        self.bearing = np.zeros((steps, batch, L), np.float32)
        p = capi.ptr
        capi.check(capi.lib().aslam_replay_generate(
            p(lm_x), p(lm_y), L, batch, steps, warm, vx, az, seed0, first_filter,
            p(self.x), p(self.y), p(self.yaw), p(self.vx), p(self.az), p(self.dt), p(self.range), p(self.bearing)))

    def odom_arrays(self):
        return self.x, self.y, self.yaw, self.vx, self.az, self.dt

    def nbytes_per_step(self):
        per = 0
        for a in (self.x, self.y, self.yaw, self.vx, self.az, self.dt, self.range, self.bearing):
            per += a.nbytes // self.steps
        return per


SEED_CFG1 = 1
SEED_CFG3 = 0x5EED0000
