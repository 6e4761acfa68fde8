"""TEST INFRASTRUCTURE ONLY -- ctypes loader of the CPU oracle (oracle/liboracle.so).

Importable from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs; never from
the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


class OrcConfig(C.Structure):
    _fields_ = [
        ("pi", C.c_float), ("min_dist_thresh", C.c_float),
        ("min_landmark_occ", C.c_int32), ("max_state_dim", C.c_int32),
        ("ukf_std_a", C.c_float), ("ukf_std_yaw", C.c_float),
        ("ukf_kp_robot_pose", C.c_float), ("ukf_kp_landmark_pose", C.c_float),
        ("ukf_kr", C.c_float), ("ukf_kq", C.c_float),
        ("ekf_kp_robot_pose", C.c_float), ("ekf_kp_landmark_pose", C.c_float),
        ("ekf_kr", C.c_float), ("ekf_kq", C.c_float),
    ]


_lib = None


def build(force=False):
    srcs = [os.path.join(_HERE, n) for n in ("oracle_c.cpp", "dense.hpp", "slam_common.hpp", "slam_filters.hpp")]
    if force or not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        l = C.CDLL(LIB_PATH)
        vp, u32, i32, f64, f32 = C.c_void_p, C.c_uint32, C.c_int32, C.c_double, C.c_float
        l.orc_config_default.argtypes = [C.POINTER(OrcConfig)]
        l.orc_create.restype = vp
        l.orc_create.argtypes = [i32, C.POINTER(OrcConfig)]
        l.orc_destroy.argtypes = [vp]
        l.orc_set_observations.argtypes = [vp, vp, vp, u32]
        l.orc_step.restype = i32
        l.orc_step.argtypes = [vp, f64, f64, f32, f64, f64, f32]
        l.orc_dim.restype = u32
        l.orc_dim.argtypes = [vp]
        l.orc_get.argtypes = [vp, vp, vp, vp]
        l.orc_get_a.argtypes = [vp, C.POINTER(f64), C.POINTER(f64)]
        l.orc_get_assoc.restype = u32
        l.orc_get_assoc.argtypes = [vp, vp, u32]
        l.orc_get_wait.restype = u32
        l.orc_get_wait.argtypes = [vp, vp, vp, vp, u32]
        l.orc_llt_fail.restype = i32
        l.orc_llt_fail.argtypes = [vp]
        l.orc_cap_rejected.restype = i32
        l.orc_cap_rejected.argtypes = [vp]
        l.orc_load_state.argtypes = [vp, u32, vp, vp, vp]
        l.orc_run_batch.restype = f64
        l.orc_run_batch.argtypes = [i32, C.POINTER(OrcConfig), u32, u32, u32, vp, vp, vp, vp, vp, vp, vp, vp,
                                    i32, u32, vp, vp, vp]
        l.orc_ekf_large_step.restype = i32
        l.orc_ekf_large_step.argtypes = [C.POINTER(OrcConfig), u32, vp, vp, C.POINTER(f64), C.POINTER(f64),
                                         f64, f64, f32, f64, f64, f32, vp, vp, u32, vp, i32]
        l.orc_num_threads.restype = i32
        _lib = l
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def default_config(**over):
    c = OrcConfig()
    lib().orc_config_default(C.byref(c))
    for k, v in over.items():
        setattr(c, k, v)
    return c


class Filter:
    """One oracle filter. kind: 'ekf' | 'ukf'."""

    def __init__(self, kind="ekf", **cfg):
        self.cfg = default_config(**cfg)
        self.kind = kind
        self._h = C.c_void_p(lib().orc_create(0 if kind == "ekf" else 1, C.byref(self.cfg)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    def set_observations(self, rng, brg):
        rng = np.ascontiguousarray(rng, np.float32)
        brg = np.ascontiguousarray(brg, np.float32)
        lib().orc_set_observations(self._h, _p(rng), _p(brg), rng.size)

    def step(self, x, y, yaw, vx, az, dt):
        return bool(lib().orc_step(self._h, x, y, np.float32(yaw), vx, az, np.float32(dt)))

    @property
    def N(self):
        return int(lib().orc_dim(self._h))

    def state(self):
        n = self.N
        X = np.zeros(n); Z = np.zeros(n); P = np.zeros((n, n))
        lib().orc_get(self._h, _p(X), _p(Z), _p(P))
        return X, Z, P.T.copy()  # column-major -> P[i, j]

    def a_entries(self):
        a, b = C.c_double(), C.c_double()
        lib().orc_get_a(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def assoc(self, cap=256):
        out = np.zeros(cap, np.int32)
        n = lib().orc_get_assoc(self._h, _p(out), cap)
        return out[:n].copy()

    def wait(self, cap=256):
        r = np.zeros(cap, np.float32); b = np.zeros(cap, np.float32); c = np.zeros(cap, np.uint32)
        n = lib().orc_get_wait(self._h, _p(r), _p(b), _p(c), cap)
        return r[:n].copy(), b[:n].copy(), c[:n].copy()

    def llt_fail(self):
        return int(lib().orc_llt_fail(self._h))

    def cap_rejected(self):
        return bool(lib().orc_cap_rejected(self._h))

    def load_state(self, X, Z, P):
        X = np.ascontiguousarray(X, np.float64); Z = np.ascontiguousarray(Z, np.float64)
        Pc = np.ascontiguousarray(np.asarray(P, np.float64).T)
        lib().orc_load_state(self._h, X.size, _p(X), _p(Z), _p(Pc))


def run_batch(kind, cfg, B, T, L, x, y, yaw, vx, az, dt, rng, brg, nthreads, ld, want_P=True):
    """B filters x T steps of a step-major replay on `nthreads` host threads.
    Returns (seconds, X [B][ld], P [B][ld][ld] indexed [b,i,j] or None, N [B])."""
    X = np.zeros((B, ld)); P = np.zeros((B, ld, ld)) if want_P else None
    N = np.zeros(B, np.uint32)
    arrs = [np.ascontiguousarray(a) for a in (x, y, yaw, vx, az, dt, rng, brg)]
    sec = lib().orc_run_batch(0 if kind == "ekf" else 1, C.byref(cfg), B, T, L, *[_p(a) for a in arrs],
                              nthreads, ld, _p(X), _p(P), _p(N))
    if P is not None:
        P = P.transpose(0, 2, 1)
    return sec, X, P, N


def ekf_large_step(cfg, X, P_colmajor, a00, a10, x, y, yaw, vx, az, dt, rng, brg, nthreads=0):
    """In-place large-map row-subset EKF step. P_colmajor: flat/2-D array holding P column-major
    (for a symmetric-to-rounding P, P[j, i] of a C array). Returns (q, assoc, a00, a10)."""
    n = X.size
    rng = np.ascontiguousarray(rng, np.float32); brg = np.ascontiguousarray(brg, np.float32)
    assoc = np.zeros(max(rng.size, 1), np.int32)
    A0, A1 = C.c_double(a00), C.c_double(a10)
    q = lib().orc_ekf_large_step(C.byref(cfg), n, _p(X), _p(P_colmajor), C.byref(A0), C.byref(A1), x, y,
                                 np.float32(yaw), vx, az, np.float32(dt), _p(rng), _p(brg), rng.size, _p(assoc),
                                 nthreads)
    return q, assoc[:rng.size], A0.value, A1.value


def laser_extract(ranges, cap=32):
    """Laser front end over scans [nscan, 360] float32 -> (range [nscan, cap], bearing [nscan, cap], counts [nscan],
    seconds). counts = -1 where the reference's assert(theta < 359) fires."""
    ranges = np.ascontiguousarray(ranges, np.float32).reshape(-1, 360)
    ns = ranges.shape[0]
    out_r = np.zeros((ns, cap), np.float32)
    out_b = np.zeros((ns, cap), np.float32)
    cnt = np.zeros(ns, np.int32)
    l = lib()
    l.orc_laser_extract.restype = C.c_double
    l.orc_laser_extract.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    sec = l.orc_laser_extract(_p(ranges), ns, cap, _p(out_r), _p(out_b), _p(cnt))
    return out_r, out_b, cnt, sec


def num_threads():
    return int(lib().orc_num_threads())
