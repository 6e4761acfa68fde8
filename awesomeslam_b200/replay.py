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
        # orientation quaternion (w, 0, 0, z) of each odometry message; yaw = quat2euler of it (tools.h:62-66)
        self.qw = np.zeros((steps, batch), np.float64)
        self.qz = np.zeros((steps, batch), np.float64)
        p = capi.ptr
        st = capi.host_lib().aslam_replay_generate_quat(
            p(lm_x), p(lm_y), L, batch, steps, warm, vx, az, seed0, first_filter,
            p(self.x), p(self.y), p(self.yaw), p(self.vx), p(self.az), p(self.dt), p(self.range), p(self.bearing),
            p(self.qw), p(self.qz))
        if st != capi.ASLAM_OK:
            raise capi.AslamError(st, "aslam_replay_generate_quat: bad arguments")

    def odom_arrays(self):
        return self.x, self.y, self.yaw, self.vx, self.az, self.dt

    def nbytes_per_step(self):
        per = 0
        for a in (self.x, self.y, self.yaw, self.vx, self.az, self.dt, self.range, self.bearing):
            per += a.nbytes // self.steps
        return per


SEED_CFG1 = 1
SEED_CFG3 = 0x5EED0000


def large_map_state(L_side, pitch=2.0, seed=4, pose=(0.0, 0.0, 0.0)):
    """Initial state of the large-map configs: landmarks on an
    L_side x L_side grid with `pitch` metres spacing centred on the origin (offset by half a pitch so
    the robot is not on a landmark); X = truth + N(0, 0.05^2); P = 0.01 I + G G' with G n x 8 ~ N(0, 0.02^2),
    pose block 1e-3 I. Returns (truth_x, truth_y, X, G, d) with P = diag(d) + G G' left implicit -- use
    dense_P() to materialise it."""
    rng = np.random.default_rng(seed)
    L = L_side * L_side
    n = 3 + 2 * L
    gx, gy = np.meshgrid(np.arange(L_side), np.arange(L_side), indexing="ij")
    tx = (gx.ravel() - (L_side - 1) / 2.0) * pitch + pitch / 2
    ty = (gy.ravel() - (L_side - 1) / 2.0) * pitch + pitch / 2
    X = np.zeros(n)
    X[0:3] = pose
    X[3::2] = tx + 0.05 * rng.standard_normal(L)
    X[4::2] = ty + 0.05 * rng.standard_normal(L)
    G = 0.02 * rng.standard_normal((n, 8))
    G[:3] = 0.0
    d = np.full(n, 0.01)
    d[:3] = 1e-3
    return tx, ty, X, G, d


def dense_P(G, d):
    P = G @ G.T
    P[np.diag_indices_from(P)] += d
    return P


def large_observations(tx, ty, pose, m=8, seed=0, sigma_r=0.01, sigma_b=0.005):
    """(range, bearing) float32 of the m landmarks nearest to `pose`, increasing bearing order."""
    rng = np.random.default_rng(seed)
    dx, dy = tx - pose[0], ty - pose[1]
    dist = np.hypot(dx, dy)
    idx = np.argsort(dist, kind="stable")[:m]
    br = np.arctan2(dy[idx], dx[idx]) - pose[2]
    br = (br + np.pi) % (2 * np.pi) - np.pi
    order = np.argsort(br, kind="stable")
    idx, br = idx[order], br[order]
    r = dist[idx] + sigma_r * rng.standard_normal(m)
    b = br + sigma_b * rng.standard_normal(m)
    return r.astype(np.float32), b.astype(np.float32), idx


# ---------------------------------------------------------------------------------------------------------------
# Synthetic LaserScan generator: the Gazebo LDS model of the reference's robot -- 360 rays, one
# per degree, 0.12 m .. 3.5 m, Gaussian range noise sigma = 0.01 m (turtlebot3_burger/model.sdf:94-119) -- against the
# map's cylinders of radius 0.25 m (models/awesome_cylinder/model.sdf:18-22). Rays without a return carry +inf, as
# Gazebo publishes them. Host-side test/bench input, plain numpy.
# ---------------------------------------------------------------------------------------------------------------
LASER_RAYS = 360
LASER_MIN, LASER_MAX, LASER_SIGMA, CYL_RADIUS = 0.12, 3.5, 0.01, 0.25


def laser_scans(lmx, lmy, poses, seed=0, sigma=LASER_SIGMA, radius=CYL_RADIUS, noise=None):
    """poses: [B, 3] (x, y, yaw) -> ranges [B, 360] float32. noise: [B, 360] standard-normal values to use instead
    of drawing them from `seed` (the CUDA generator aslam_laser_synthesize takes the same array)."""
    poses = np.asarray(poses, np.float64).reshape(-1, 3)
    B = poses.shape[0]
    lm = np.stack([np.asarray(lmx, np.float64), np.asarray(lmy, np.float64)], axis=1)  # [L, 2]
    ang = poses[:, 2:3] + np.deg2rad(np.arange(LASER_RAYS, dtype=np.float64))[None, :]  # [B, 360]
    dx, dy = np.cos(ang), np.sin(ang)
    best = np.full((B, LASER_RAYS), np.inf)
    for l in range(lm.shape[0]):  # ray-circle intersection, nearest positive root
        cx = lm[l, 0] - poses[:, 0:1]
        cy = lm[l, 1] - poses[:, 1:2]
        b = dx * cx + dy * cy
        disc = b * b - (cx * cx + cy * cy - radius * radius)
        with np.errstate(invalid="ignore"):
            t = b - np.sqrt(disc)
        hit = (disc >= 0) & (t > 0)
        best = np.where(hit & (t < best), t, best)
    if noise is None:
        noise = np.random.default_rng(seed).standard_normal(best.shape)
    noisy = best + sigma * np.asarray(noise, np.float64).reshape(best.shape)
    out = np.where((best >= LASER_MIN) & (best <= LASER_MAX), noisy, np.inf)
    return out.astype(np.float32)


LASER_POSES_CFG3 = dict(spread=0.2, yaw_deg=(19.0, 26.0))  # ray 0 between the 0 and 45 degree cylinders


def laser_poses(B, seed=0, spread=0.6, yaw_deg=(20.0, 70.0)):
    """Robot poses around the origin whose ray 0 / ray 359 do not graze a cylinder (the reference aborts when the
    first and the last ray fall on the same object, sensor_landmark.cpp:82). Defaults suit the cfg1 map (cylinders
    at 0, 90, 180, 270 degrees and on the diagonals at 4.2 m, out of range); for the cfg3 map, which also has
    cylinders on the diagonals at 2.1 m, use LASER_POSES_CFG3."""
    rng = np.random.default_rng(seed)
    x = spread * (rng.random(B) - 0.5)
    y = spread * (rng.random(B) - 0.5)
    yaw = np.deg2rad(yaw_deg[0] + (yaw_deg[1] - yaw_deg[0]) * rng.random(B))
    return np.stack([x, y, yaw], axis=1)
