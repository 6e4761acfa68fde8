"""Headless stand-ins for the two ends of the reference's ROS graph. Host-side glue, no GPU.

* `marker_list` / `write_markers_csv` / `write_markers_json`: what `aslam::VisualizeLandmark::callback`
  (awesome_slam/src/visualize_landmark/visualize_landmark.cpp:47-72, `createMarker` :74-99) would publish as a
  MarkerArray for RViz -- ground-truth cylinders (blue) followed by the filter's estimates (green), ids counted
  through both lists, positions rounded to float as `createMarker(const float&, const float&, ...)` does -- written to
  a file instead (no graphics stack on this box).
* `save_replay` / `load_replay`: a rosbag-free binary replay of the two topics the filter nodes subscribe to
  (`/odom` fields read at src/ekf/ekf.cpp:94,139-142 and `out/landmarks/sensor`, awesome_slam_msgs/Landmarks), in the
  step-major layout `aslam_filter_run_host` takes. Format (little endian):
      magic  8 bytes  b"ASLAMRP1"
      u32 batch, u32 steps, u32 obs_per_filter, u32 reserved (0)
      f64 x[steps][batch], f64 y[steps][batch], f64 vx[steps][batch], f64 az[steps][batch]
      f32 yaw[steps][batch], f32 dt[steps][batch]
      f32 range[steps][batch][obs_per_filter], f32 bearing[steps][batch][obs_per_filter]
"""
import csv
import json
import struct

import numpy as np

MARKER_FIELDS = ("id", "kind", "x", "y", "z", "scale_x", "scale_y", "scale_z", "r", "g", "b", "a", "frame_id", "type")
REPLAY_MAGIC = b"ASLAMRP1"


def _marker(pos_x, pos_y, green, blue, mid, kind):
    # createMarker (visualize_landmark.cpp:74-99): positions arrive as `const float&`
    return {"id": mid, "kind": kind, "x": float(np.float32(pos_x)), "y": float(np.float32(pos_y)), "z": 0.0,
            "scale_x": 0.5, "scale_y": 0.5, "scale_z": 1.0, "r": 0.0, "g": float(green), "b": float(blue), "a": 1.0,
            "frame_id": "odom", "type": "CYLINDER"}


def marker_list(orig_x, orig_y, est_x, est_y):
    """visualize_landmark.cpp:47-72: ground truth first (g = 0, b = 1), then the estimates (g = 1, b = 0); the id runs
    through both (`++id` from -1)."""
    out, mid = [], -1
    for x, y in zip(orig_x, orig_y):
        mid += 1
        out.append(_marker(x, y, 0.0, 1.0, mid, "truth"))
    for x, y in zip(est_x, est_y):
        mid += 1
        out.append(_marker(x, y, 1.0, 0.0, mid, "estimate"))
    return out


def write_markers_csv(path, markers):
    with open(path, "w", newline="") as fh:
        w = csv.DictWriter(fh, fieldnames=MARKER_FIELDS)
        w.writeheader()
        for m in markers:
            w.writerow(m)


def write_markers_json(path, markers):
    with open(path, "w") as fh:
        json.dump({"markers": markers}, fh, indent=1)


def save_replay(path, x, y, yaw, vx, az, dt, rng, brg):
    """Arrays as `Replay` holds them: x, y, vx, az float64 [steps, batch]; yaw, dt float32 [steps, batch];
    rng, brg float32 [steps, batch, obs]."""
    x = np.ascontiguousarray(x, np.float64)
    steps, batch = x.shape
    rng = np.ascontiguousarray(rng, np.float32).reshape(steps, batch, -1)
    obs = rng.shape[2]
    with open(path, "wb") as fh:
        fh.write(REPLAY_MAGIC)
        fh.write(struct.pack("<4I", batch, steps, obs, 0))
        for a in (x, y, vx, az):
            fh.write(np.ascontiguousarray(a, "<f8").reshape(steps, batch).tobytes())
        for a in (yaw, dt):
            fh.write(np.ascontiguousarray(a, "<f4").reshape(steps, batch).tobytes())
        fh.write(rng.astype("<f4").tobytes())
        fh.write(np.ascontiguousarray(brg, "<f4").reshape(steps, batch, obs).tobytes())


def load_replay(path):
    """-> dict(x, y, vx, az, yaw, dt, range, bearing, batch, steps, obs_per_filter)."""
    with open(path, "rb") as fh:
        if fh.read(8) != REPLAY_MAGIC:
            raise ValueError("not an ASLAMRP1 replay file")
        batch, steps, obs, _ = struct.unpack("<4I", fh.read(16))
        n = steps * batch

        def rd(dtype, count, shape):
            a = np.frombuffer(fh.read(count * np.dtype(dtype).itemsize), dtype=dtype, count=count)
            if a.size != count:
                raise ValueError("truncated replay file")
            return a.reshape(shape).copy()

        out = {"batch": batch, "steps": steps, "obs_per_filter": obs}
        for k in ("x", "y", "vx", "az"):
            out[k] = rd("<f8", n, (steps, batch))
        for k in ("yaw", "dt"):
            out[k] = rd("<f4", n, (steps, batch))
        out["range"] = rd("<f4", n * obs, (steps, batch, obs))
        out["bearing"] = rd("<f4", n * obs, (steps, batch, obs))
    return out
