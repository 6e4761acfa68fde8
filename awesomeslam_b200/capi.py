"""ctypes binding of libaslam_b200.so (include/aslam_c.h).

This is the same stub a maintainer of the reference would add on their side of the boundary
(INTEGRATION.md shows the C++ one). There is NO CPU path behind it: if the CUDA library has
not been built, `lib()` raises; if no CUDA device is usable, the create calls return
ASLAM_ERR_CUDA and `check` raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libaslam_b200.so")

ASLAM_OK = 0
ASLAM_ERR_ARG = -1
ASLAM_ERR_CUDA = -2
ASLAM_ERR_UNSUPPORTED = -3
ASLAM_ERR_NOT_READY = -4
ASLAM_ERR_TIMEOUT = -5

ST_CAP_REJECTED = 1 << 0
ST_LLT_FAILED = 1 << 1
ST_NONFINITE = 1 << 2
ST_WAIT_OVERFLOW = 1 << 3
ST_OBS_OVERFLOW = 1 << 4

MAX_WAIT = 256
MAX_OBS = 64
KIND_EKF = 0
KIND_UKF = 1


class AslamError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("aslam status %d: %s" % (status, msg))
        self.status = status


class Config(C.Structure):
    """aslam_config == include/awesome_slam/config.h:39-65 of the reference."""
    _fields_ = [
        ("pi", C.c_float), ("min_dist_thresh", C.c_float),
        ("min_landmark_occ", C.c_int32), ("max_state_dim", C.c_int32),
        ("ukf_std_a", C.c_float), ("ukf_std_yaw", C.c_float),
        ("ukf_kp_robot_pose", C.c_float), ("ukf_kp_landmark_pose", C.c_float),
        ("ukf_kr", C.c_float), ("ukf_kq", C.c_float),
        ("ekf_kp_robot_pose", C.c_float), ("ekf_kp_landmark_pose", C.c_float),
        ("ekf_kr", C.c_float), ("ekf_kq", C.c_float),
    ]


class Odom(C.Structure):
    """aslam_odom: SoA pointers (host or device, depending on the call)."""
    _fields_ = [("x", C.c_void_p), ("y", C.c_void_p), ("yaw", C.c_void_p),
                ("vx", C.c_void_p), ("az", C.c_void_p), ("dt", C.c_void_p)]


_vp = C.c_void_p
_u32 = C.c_uint32
_i32 = C.c_int32
_f64 = C.c_double
_f32 = C.c_float
_sz = C.c_size_t

# name -> (restype, argtypes); every symbol include/aslam_c.h declares
SIGNATURES = {
    "aslam_config_default": (None, [C.POINTER(Config)]),
    "aslam_last_error": (C.c_char_p, []),
    "aslam_version": (C.c_char_p, []),
    "aslam_device_count": (_i32, []),
    "aslam_quat2euler": (_f32, [_f64, _f64, _f64, _f64]),
    "aslam_probe_fp64_tflops": (_i32, [_i32, C.POINTER(_f64)]),
    "aslam_filter_read_state": (_i32, [_vp, _vp]),
    "aslam_filter_read_state_async": (_i32, [_vp, _vp, C.POINTER(C.c_uint64)]),
    "aslam_filter_read_wait": (_i32, [_vp, C.c_uint64]),
    "aslam_ekf_create": (_i32, [C.POINTER(Config), _u32, _i32, C.POINTER(_vp)]),
    "aslam_ukf_create": (_i32, [C.POINTER(Config), _u32, _i32, C.POINTER(_vp)]),
    "aslam_filter_destroy": (_i32, [_vp]),
    "aslam_filter_sync": (_i32, [_vp]),
    "aslam_filter_stream": (_vp, [_vp]),
    "aslam_filter_batch": (_u32, [_vp]),
    "aslam_filter_ld": (_u32, [_vp]),
    "aslam_filter_kind": (_i32, [_vp]),
    "aslam_filter_set_observations": (_i32, [_vp, _vp, _vp, _vp]),
    "aslam_filter_set_observations_device": (_i32, [_vp, _vp, _vp, _vp]),
    "aslam_filter_step": (_i32, [_vp, C.POINTER(Odom)]),
    "aslam_filter_step_device": (_i32, [_vp, C.POINTER(Odom)]),
    "aslam_filter_run_device": (_i32, [_vp, _u32, C.POINTER(Odom), _vp, _vp, _u32]),
    "aslam_filter_run_host": (_i32, [_vp, _u32, C.POINTER(Odom), _vp, _vp, _u32]),
    "aslam_filter_get_state": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "aslam_filter_get_aux": (_i32, [_vp, _vp, _vp, _vp]),
    "aslam_filter_get_wait": (_i32, [_vp, _vp, _vp, _vp, _vp]),
    "aslam_filter_get_landmarks": (_i32, [_vp, _vp, _vp, _vp]),
    "aslam_filter_get_llt_fail": (_i32, [_vp, _vp]),
    "aslam_filter_device_state": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "aslam_filter_load_state": (_i32, [_vp, _vp, _vp, _vp, _vp]),
    "aslam_filter_load_aux": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "aslam_filter_launch_count": (C.c_uint64, [_vp]),
    "aslam_large_ekf_create": (_i32, [C.POINTER(Config), _u32, _u32, _u32, _i32, C.POINTER(_vp)]),
    "aslam_large_ukf_create": (_i32, [C.POINTER(Config), _u32, _i32, C.POINTER(_vp)]),
    "aslam_large_destroy": (_i32, [_vp]),
    "aslam_large_sync": (_i32, [_vp]),
    "aslam_large_stream": (_vp, [_vp]),
    "aslam_large_load_state": (_i32, [_vp, _vp, _vp, _vp, _sz]),
    "aslam_large_get_state": (_i32, [_vp, _vp, _vp, _sz]),
    "aslam_large_step": (_i32, [_vp, _f64, _f64, _f32, _f64, _f64, _f32, _vp, _vp, _u32, _vp]),
    "aslam_large_step_phase1": (_i32, [_vp, _f64, _f64, _f32, _f64, _f64, _f32, _vp, _vp, _u32, C.POINTER(_u32)]),
    "aslam_large_exchange_buffer": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "aslam_large_step_phase2": (_i32, [_vp, _vp]),
    "aslam_laser_config_default": (None, [_vp]),
    "aslam_laser_create": (_i32, [_vp, _u32, _i32, C.POINTER(_vp)]),
    "aslam_laser_destroy": (_i32, [_vp]),
    "aslam_laser_extract": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp]),
    "aslam_laser_extract_device": (_i32, [_vp, _vp, _u32, _vp, _vp, _vp]),
    "aslam_laser_synthesize": (_i32, [_vp, _vp, _vp, _u32, _vp, _vp, _u32, _f64, _f64, _vp]),
    "aslam_laser_synthesize_device": (_i32, [_vp, _vp, _vp, _u32, _vp, _vp, _u32, _f64, _f64, _vp]),
    "aslam_laser_sync": (_i32, [_vp]),
    "aslam_laser_stream": (_vp, [_vp]),
    "aslam_laser_launch_count": (C.c_uint64, [_vp]),
    "aslam_large_peer_export": (_i32, [_vp, _vp, C.POINTER(_vp)]),
    "aslam_large_peer_attach": (_i32, [_vp, _u32, _u32, _vp, C.POINTER(_vp)]),
    "aslam_large_step_fused": (_i32, [_vp, _f64, _f64, _f32, _f64, _f64, _f32, _vp, _vp, _u32, _vp]),
    "aslam_large_launch_count": (C.c_uint64, [_vp]),
    "aslam_large_set_timing": (_i32, [_vp, _i32]),
    "aslam_large_trace": (_i32, [_vp, _vp]),
    "aslam_large_last_kernel_ms": (_i32, [_vp, C.POINTER(_f32)]),
    "aslam_large_status": (_i32, [_vp, C.POINTER(_u32), C.POINTER(_i32)]),
    "aslam_dense_gemm_nt": (_i32, [_vp, _i32, _i32, _i32, _vp, _sz, _vp, _sz, _vp, _sz, _f64, _f64, _i32]),
    "aslam_dense_cholesky": (_i32, [_vp, _vp, _sz, _i32, _vp, _vp]),
    "aslam_dense_trsm": (_i32, [_vp, _vp, _sz, _i32, _vp, _sz, _i32, _vp]),
    "aslam_probe_dmma_tflops": (_i32, [_i32, C.POINTER(_f64)]),
    "aslam_large_ukf_set_precision": (_i32, [_vp, _i32]),
    "aslam_large_ukf_products_ms": (_i32, [_vp, C.POINTER(C.c_float)]),
    "aslam_dense_gemm_nt_tf32": (_i32, [_vp, _i32, _i32, _i32, _vp, _sz, _vp, _sz, _vp, _sz, _f64, _f64, _i32, _i32]),
    "aslam_dense_gemm_nt_tf32_static": (_i32, [_vp, _i32, _i32, _i32, _vp, _sz, _vp, _sz, _vp, _sz, _f64, _f64, _i32, _i32]),
    "aslam_probe_tf32_tflops": (_i32, [_i32, C.POINTER(_f64)]),
    "aslam_replay_generate": (_i32, [_vp, _vp, _u32, _u32, _u32, _u32, _f64, _f64, C.c_uint64, _u32,
                                     _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "aslam_replay_generate_quat": (_i32, [_vp, _vp, _u32, _u32, _u32, _u32, _f64, _f64, C.c_uint64, _u32,
                                          _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
}

# The host-only helpers (no CUDA behind them): csrc/replay.cpp built alone into lib/libaslam_host.so, so that input
# generation -- e.g. for bench.py's reference arm -- does not have to map the CUDA library.
HOST_LIB_PATH = os.path.join(_HERE, "lib", "libaslam_host.so")
HOST_SYMBOLS = ("aslam_quat2euler", "aslam_replay_generate", "aslam_replay_generate_quat")
_host = None


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise ImportError("%s is missing: build it with `python -m awesomeslam_b200.build`" % HOST_LIB_PATH)
        l = C.CDLL(HOST_LIB_PATH)
        for name in HOST_SYMBOLS:
            fn = getattr(l, name)
            fn.restype, fn.argtypes = SIGNATURES[name]
        _host = l
    return _host

_lib = None


def lib():
    """Loads the CUDA library; raises (never falls back) when it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("ASLAM_B200_LIB", LIB_PATH)  # experiments: another build of the same library
    if not os.path.exists(path):
        raise ImportError(
            "%s is missing: build it with `python -m awesomeslam_b200.build` "
            "(there is no CPU implementation behind this package)" % path)
    l = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(l, name)
        fn.restype = res
        fn.argtypes = args
    _lib = l
    return l


def check(status):
    if status != ASLAM_OK:
        raise AslamError(status, lib().aslam_last_error().decode("utf-8", "replace"))


def default_config(**overrides):
    cfg = Config()
    lib().aslam_config_default(C.byref(cfg))
    for k, v in overrides.items():
        if not hasattr(cfg, k):
            raise AttributeError(k)
        setattr(cfg, k, v)
    return cfg


def ptr(a):
    """void* of a numpy array (must be C-contiguous) or None."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def as_np(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)
