"""TEST INFRASTRUCTURE ONLY -- ctypes loader of oracle/_ref/: the reference's OWN ekf.cpp / ukf.cpp /
sensor_landmark.cpp, compiled unmodified against the Eigen / ROS stand-ins of oracle/ref_shim (recipe:
oracle/ref_shim/Makefile). Used to pin the restated oracle (tests/test_ref_pin.py) and as the CPU baseline of
bench.py (cpu_baseline.kind = "reference"). Never imported by the product package.

Each library is loaded RTLD_LOCAL: ekf.cpp and ukf.cpp both define a global `struct Parameters` and the non-inline
functions of common.h, so they must not share a symbol namespace.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
REF_SRC = "/root/reference/awesome_slam"
_libs = {}


def available(kind="ekf", cap64=False):
    return os.path.exists(_path(kind, cap64)) or os.path.isdir(REF_SRC)


def _path(kind, cap64):
    return os.path.join(REF_DIR, "libref_%s%s.so" % (kind, "_cap64" if cap64 else ""))


def build(force=False):
    """Runs the committed recipe when the reference sources are on disk (this container); otherwise keeps the
    prebuilt libraries that travelled with the snapshot."""
    if os.path.isdir(REF_SRC):
        subprocess.check_call(["make", "-C", os.path.join(_HERE, "ref_shim"), "REF=" + REF_SRC] +
                              (["-B"] if force else []), stdout=subprocess.DEVNULL)


def lib(kind="ekf", cap64=False):
    key = (kind, cap64)
    if key not in _libs:
        if not os.path.exists(_path(kind, cap64)):
            build()
        l = C.CDLL(_path(kind, cap64), mode=os.RTLD_LOCAL)
        vp, u32, i32, f64, f32 = C.c_void_p, C.c_uint32, C.c_int32, C.c_double, C.c_float
        if kind in ("ekf", "ukf"):
            l.ref_create.restype = vp
            l.ref_destroy.argtypes = [vp]
            l.ref_set_time.argtypes = [vp, f64]
            l.ref_landmarks.argtypes = [vp, vp, vp, u32]
            l.ref_odom.argtypes = [vp, f64, f64, f64, f64, f64, f64, f64, f64]
            l.ref_dim.restype = u32
            l.ref_dim.argtypes = [vp]
            l.ref_flags.restype = i32
            l.ref_flags.argtypes = [vp]
            l.ref_get.argtypes = [vp, vp, vp, vp]
            l.ref_get_a.restype = i32
            l.ref_get_a.argtypes = [vp, C.POINTER(f64), C.POINTER(f64)]
            l.ref_get_wait.restype = u32
            l.ref_get_wait.argtypes = [vp, vp, vp, vp, u32]
            l.ref_get_sensor.restype = u32
            l.ref_get_sensor.argtypes = [vp, vp, vp, u32]
            l.ref_get_published.restype = C.c_uint64
            l.ref_get_published.argtypes = [vp, vp, vp, u32, C.POINTER(u32)]
            l.ref_quat2euler.restype = f32
            l.ref_quat2euler.argtypes = [f32, f32, f32, f32]
            l.ref_normalize_angle.restype = f32
            l.ref_normalize_angle.argtypes = [f32]
            l.ref_max_state_dim.restype = i32
            l.ref_run_batch.restype = f64
            l.ref_run_batch.argtypes = [u32, u32, u32, vp, vp, vp, vp, vp, vp, vp, vp, i32, u32, vp, vp, vp]
            l.ref_num_threads.restype = i32
        else:
            l.ref_laser_create.restype = vp
            l.ref_laser_destroy.argtypes = [vp]
            l.ref_laser_scan.restype = i32
            l.ref_laser_scan.argtypes = [vp, vp, vp, vp, i32]
            l.ref_laser_extract.restype = f64
            l.ref_laser_extract.argtypes = [vp, i32, i32, vp, vp, vp, i32]
        _libs[key] = l
    return _libs[key]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Node:
    """One reference node object (aslam::EKFSlam / aslam::UKFSlam) driven through its ROS callbacks."""

    def __init__(self, kind="ekf", cap64=False):
        self.kind = kind
        self.l = lib(kind, cap64)
        self._h = C.c_void_p(self.l.ref_create())

    def __del__(self):
        if getattr(self, "_h", None):
            self.l.ref_destroy(self._h)
            self._h = None

    def set_time(self, t):
        self.l.ref_set_time(self._h, float(t))

    def landmarks(self, x, y):
        """cbSensorLandmark with a Landmarks message (float64 x = range, y = bearing)."""
        x = np.ascontiguousarray(x, np.float64)
        y = np.ascontiguousarray(y, np.float64)
        self.l.ref_landmarks(self._h, _p(x), _p(y), x.size)

    def odom(self, px, py, qw, qx, qy, qz, vx, az):
        """cbOdom with a nav_msgs/Odometry message."""
        self.l.ref_odom(self._h, px, py, qw, qx, qy, qz, vx, az)

    @property
    def N(self):
        return int(self.l.ref_dim(self._h))

    @property
    def flags(self):
        f = self.l.ref_flags(self._h)
        return bool(f & 1), bool(f & 2)  # init_x, init_z

    def state(self):
        n = self.N
        X = np.zeros(n); Z = np.zeros(n); P = np.zeros((n, n))
        self.l.ref_get(self._h, _p(X), _p(Z), _p(P))
        return X, Z, P.T.copy()

    def a_entries(self):
        a, b = C.c_double(), C.c_double()
        return (a.value, b.value) if self.l.ref_get_a(self._h, C.byref(a), C.byref(b)) else None

    def wait(self, cap=1024):
        r = np.zeros(cap, np.float32); b = np.zeros(cap, np.float32); c = np.zeros(cap, np.uint32)
        n = self.l.ref_get_wait(self._h, _p(r), _p(b), _p(c), cap)
        return r[:n].copy(), b[:n].copy(), c[:n].copy()

    def sensor(self, cap=1024):
        r = np.zeros(cap, np.float32); b = np.zeros(cap, np.float32)
        n = self.l.ref_get_sensor(self._h, _p(r), _p(b), cap)
        return r[:n].copy(), b[:n].copy()

    def published(self, cap=1024):
        """(x, y, number of messages published so far) of out/landmarks/kalman."""
        x = np.zeros(cap); y = np.zeros(cap)
        cnt = C.c_uint32()
        total = self.l.ref_get_published(self._h, _p(x), _p(y), cap, C.byref(cnt))
        return x[:cnt.value].copy(), y[:cnt.value].copy(), int(total)


def quat2euler(w, x, y, z, kind="ekf"):
    return float(lib(kind).ref_quat2euler(w, x, y, z))


def run_batch(kind, B, T, L, x, y, qw, qz, vx, az, rng, brg, nthreads, ld, want_P=True, cap64=False):
    """B reference nodes x T messages of a step-major replay (dt = 1.0f) on `nthreads` host threads.
    Returns (seconds, X [B][ld], P [B][ld][ld] indexed [b,i,j] or None, N [B])."""
    X = np.zeros((B, ld)); P = np.zeros((B, ld, ld)) if want_P else None
    N = np.zeros(B, np.uint32)
    arrs = [np.ascontiguousarray(a) for a in (x, y, qw, qz, vx, az, rng, brg)]
    assert arrs[6].dtype == np.float32 and arrs[7].dtype == np.float32 and arrs[0].dtype == np.float64
    sec = lib(kind, cap64).ref_run_batch(B, T, L, *[_p(a) for a in arrs], nthreads, ld, _p(X), _p(P), _p(N))
    if P is not None:
        P = P.transpose(0, 2, 1)
    return sec, X, P, N


def num_threads(kind="ekf"):
    return int(lib(kind).ref_num_threads())


def laser_extract(ranges, cap=32, nthreads=0):
    """The reference's LandMarks::callback over scans [nscan, 360] float32 -> (range [nscan, cap], bearing, counts,
    seconds); counts = -1 where its assert(theta < 359) would fire (the driver checks the same condition first)."""
    ranges = np.ascontiguousarray(ranges, np.float32).reshape(-1, 360)
    ns = ranges.shape[0]
    out_r = np.zeros((ns, cap), np.float32); out_b = np.zeros((ns, cap), np.float32)
    cnt = np.zeros(ns, np.int32)
    sec = lib("laser").ref_laser_extract(_p(ranges), ns, cap, _p(out_r), _p(out_b), _p(cnt), nthreads)
    return out_r, out_b, cnt, sec
