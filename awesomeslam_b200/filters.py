"""Host-side mirror of the reference's filter classes over the C ABI.

The reference's "API" is two ROS node classes, aslam::EKFSlam (src/ekf/ekf.h:71-132) and
aslam::UKFSlam (src/ukf/ukf.h:86-144), whose callbacks are

    cbSensorLandmark(Landmarks{x=range[], y=bearing[]})      ekf.cpp:102-114, ukf.cpp:98-110
    cbOdom(Odometry)  -> publishes Landmarks{x[], y[]}        ekf.cpp:74-99,   ukf.cpp:70-95

`EKFSlam` / `UKFSlam` below keep those names and argument meanings for ONE filter (batch = 1);
`BatchedFilter` is the same thing for B independent filters (structure-of-arrays messages).
All arithmetic happens in the CUDA library; this file only marshals buffers.
"""
import ctypes as C

import numpy as np

from . import capi


class Odometry:
    """The nav_msgs/Odometry fields the reference reads (ekf.cpp:94,139-142,206-209)."""

    __slots__ = ("x", "y", "qw", "qx", "qy", "qz", "vx", "az")

    def __init__(self, x=0.0, y=0.0, qw=1.0, qx=0.0, qy=0.0, qz=0.0, vx=0.0, az=0.0):
        self.x, self.y = x, y
        self.qw, self.qx, self.qy, self.qz = qw, qx, qy, qz
        self.vx, self.az = vx, az


class BatchedFilter:
    """B independent EKF- or UKF-SLAM filters on one GPU (aslam_filter handle)."""

    def __init__(self, kind="ekf", batch=1, device=0, config=None, **cfg_overrides):
        self._h = C.c_void_p()
        self._lib = capi.lib()
        self.config = config if config is not None else capi.default_config(**cfg_overrides)
        self.kind = kind
        create = {"ekf": self._lib.aslam_ekf_create, "ukf": self._lib.aslam_ukf_create}[kind]
        capi.check(create(C.byref(self.config), batch, device, C.byref(self._h)))
        self.batch = batch
        self.device = device
        self.ld = self._lib.aslam_filter_ld(self._h)
        self._keep = []  # host arrays referenced by in-flight async copies
        self._obs_total = 0

    # -- lifetime -------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.aslam_filter_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    @property
    def stream(self):
        return self._lib.aslam_filter_stream(self._h)

    def sync(self):
        capi.check(self._lib.aslam_filter_sync(self._h))
        self._keep.clear()

    # -- cbSensorLandmark -----------------------------------------------------------------
    def set_observations(self, rng, brg, offsets=None):
        """rng/brg: float32 [total]; offsets: uint32 [batch+1] (None: equal split)."""
        rng = capi.as_np(rng, np.float32).ravel()
        brg = capi.as_np(brg, np.float32).ravel()
        if offsets is None:
            per = rng.size // self.batch
            assert per * self.batch == rng.size
            offsets = np.arange(self.batch + 1, dtype=np.uint32) * np.uint32(per)
        offsets = capi.as_np(offsets, np.uint32)
        assert offsets.size == self.batch + 1 and int(offsets[-1]) == rng.size == brg.size
        self._keep += [rng, brg, offsets]
        self._obs_total = int(rng.size)
        capi.check(self._lib.aslam_filter_set_observations(self._h, capi.ptr(rng), capi.ptr(brg), capi.ptr(offsets)))

    # -- cbOdom ---------------------------------------------------------------------------
    def step(self, x, y, yaw, vx, az, dt):
        """One odometry message per filter. yaw: float32 quat2euler result; dt: float32."""
        arrs = [capi.as_np(x, np.float64).ravel(), capi.as_np(y, np.float64).ravel(),
                capi.as_np(yaw, np.float32).ravel(), capi.as_np(vx, np.float64).ravel(),
                capi.as_np(az, np.float64).ravel(), capi.as_np(dt, np.float32).ravel()]
        for a in arrs:
            assert a.size == self.batch
        self._keep += arrs
        od = capi.Odom(*[capi.ptr(a) for a in arrs])
        capi.check(self._lib.aslam_filter_step(self._h, C.byref(od)))

    def run_host(self, steps, x, y, yaw, vx, az, dt, rng, brg, obs_per_filter):
        """`steps` messages from HOST buffers (step-major [steps][batch](...[obs])): copies them to
        the device and replays them in one launch. Asynchronous; call sync()/get_state()."""
        arrs = [capi.as_np(x, np.float64), capi.as_np(y, np.float64), capi.as_np(yaw, np.float32),
                capi.as_np(vx, np.float64), capi.as_np(az, np.float64), capi.as_np(dt, np.float32)]
        rng = capi.as_np(rng, np.float32)
        brg = capi.as_np(brg, np.float32)
        for a in arrs:
            assert a.size == steps * self.batch
        assert rng.size == brg.size == steps * self.batch * obs_per_filter
        self._keep += arrs + [rng, brg]
        od = capi.Odom(*[capi.ptr(a) for a in arrs])
        self._obs_total = self.batch * obs_per_filter
        capi.check(self._lib.aslam_filter_run_host(self._h, steps, C.byref(od), capi.ptr(rng), capi.ptr(brg),
                                                   obs_per_filter))

    def run_device(self, steps, d_x, d_y, d_yaw, d_vx, d_az, d_dt, d_rng, d_brg, obs_per_filter):
        """Same with device pointers (ints), e.g. torch tensors' data_ptr(). No copies."""
        od = capi.Odom(d_x, d_y, d_yaw, d_vx, d_az, d_dt)
        self._obs_total = self.batch * obs_per_filter
        capi.check(self._lib.aslam_filter_run_device(self._h, steps, C.byref(od), d_rng, d_brg, obs_per_filter))

    # -- state export ---------------------------------------------------------------------
    def get_state(self, want_P=True, want_assoc=True):
        B, ld = self.batch, self.ld
        X = np.zeros((B, ld), np.float64)
        P = np.zeros((B, ld, ld), np.float64) if want_P else None
        N = np.zeros(B, np.uint32)
        st = np.zeros(B, np.uint32)
        assoc = np.full(max(self._obs_total, 1), -2, np.int32) if want_assoc else None
        capi.check(self._lib.aslam_filter_get_state(self._h, capi.ptr(X), capi.ptr(P), capi.ptr(N),
                                                    capi.ptr(assoc), capi.ptr(st)))
        self._keep.clear()
        if P is not None:
            P = P.transpose(0, 2, 1)  # exported column-major -> index as P[b, i, j]
        if assoc is not None:
            assoc = assoc[:self._obs_total]
        return {"X": X, "P": P, "N": N, "status": st, "assoc": assoc}

    def read_state_into(self, X_out):
        """param.X of every filter into a caller array [batch, ld] (may be pinned). Synchronises."""
        assert X_out.dtype == np.float64 and X_out.size == self.batch * self.ld and X_out.flags["C_CONTIGUOUS"]
        capi.check(self._lib.aslam_filter_read_state(self._h, capi.ptr(X_out)))
        self._keep.clear()

    def read_state_async(self, X_out):
        """Pipelined read of param.X: returns a ticket at once; X_out (keep it alive, ideally pinned) is
        complete after read_wait(ticket) or sync(). Does not stall the next run_host/run_device call."""
        assert X_out.dtype == np.float64 and X_out.size == self.batch * self.ld and X_out.flags["C_CONTIGUOUS"]
        t = C.c_uint64(0)
        capi.check(self._lib.aslam_filter_read_state_async(self._h, capi.ptr(X_out), C.byref(t)))
        return int(t.value)

    def read_wait(self, ticket):
        capi.check(self._lib.aslam_filter_read_wait(self._h, ticket))

    def get_aux(self):
        B, ld = self.batch, self.ld
        Z = np.zeros((B, ld), np.float64)
        a00 = np.zeros(B, np.float64)
        a10 = np.zeros(B, np.float64)
        capi.check(self._lib.aslam_filter_get_aux(self._h, capi.ptr(Z), capi.ptr(a00), capi.ptr(a10)))
        return {"Z": Z, "a00": a00, "a10": a10}

    def get_wait(self):
        B, W = self.batch, capi.MAX_WAIT
        cnt = np.zeros(B, np.uint32)
        r = np.zeros((B, W), np.float32)
        b = np.zeros((B, W), np.float32)
        h = np.zeros((B, W), np.uint32)
        capi.check(self._lib.aslam_filter_get_wait(self._h, capi.ptr(cnt), capi.ptr(r), capi.ptr(b), capi.ptr(h)))
        return {"count": cnt, "range": r, "bearing": b, "hits": h}

    def get_llt_fail(self):
        out = np.zeros(self.batch, np.int32)
        capi.check(self._lib.aslam_filter_get_llt_fail(self._h, capi.ptr(out)))
        return out

    def get_landmarks(self):
        """convertToLandmarkMsg (common.h:93-108): per filter the mapped landmark x[], y[]."""
        B, lmax = self.batch, (self.ld - 3) // 2
        x = np.zeros((B, lmax), np.float64)
        y = np.zeros((B, lmax), np.float64)
        c = np.zeros(B, np.uint32)
        capi.check(self._lib.aslam_filter_get_landmarks(self._h, capi.ptr(x), capi.ptr(y), capi.ptr(c)))
        return x, y, c

    def load_state(self, N, X, Z, P):
        """P given as [B, i, j]; padded to ld."""
        B, ld = self.batch, self.ld
        N = capi.as_np(N, np.uint32)
        Xp = np.zeros((B, ld)); Zp = np.zeros((B, ld)); Pp = np.zeros((B, ld, ld))
        for b in range(B):
            n = int(N[b])
            Xp[b, :n] = X[b][:n]
            Zp[b, :n] = Z[b][:n]
            Pp[b, :n, :n] = np.asarray(P[b])[:n, :n].T  # column-major on the wire
        capi.check(self._lib.aslam_filter_load_state(self._h, capi.ptr(N), capi.ptr(Xp), capi.ptr(Zp), capi.ptr(Pp)))

    def load_aux(self, a00=None, a10=None, wait=None, status=None):
        """The rest of a checkpoint (after load_state): A entries from get_aux, wait = get_wait()'s dict, status."""
        a00 = None if a00 is None else capi.as_np(a00, np.float64)
        a10 = None if a10 is None else capi.as_np(a10, np.float64)
        wc = wr = wb = wh = None
        if wait is not None:
            wc = capi.as_np(wait["count"], np.uint32)
            wr = capi.as_np(wait["range"], np.float32)
            wb = capi.as_np(wait["bearing"], np.float32)
            wh = capi.as_np(wait["hits"], np.uint32)
        st = None if status is None else capi.as_np(status, np.uint32)
        capi.check(self._lib.aslam_filter_load_aux(self._h, capi.ptr(a00), capi.ptr(a10), capi.ptr(wc), capi.ptr(wr),
                                                   capi.ptr(wb), capi.ptr(wh), capi.ptr(st)))

    @property
    def launch_count(self):
        return int(self._lib.aslam_filter_launch_count(self._h))


class _SingleFilter:
    """One filter with the reference's callback names."""

    KIND = "ekf"

    def __init__(self, device=0, config=None, **cfg_overrides):
        self._f = BatchedFilter(self.KIND, 1, device, config, **cfg_overrides)
        self._last_time = None

    def close(self):
        self._f.close()

    def cbSensorLandmark(self, x, y):
        """msg.x = ranges, msg.y = bearings (float64 on the wire, stored as float: ekf.cpp:111)."""
        self._f.set_observations(np.asarray(x, np.float64).astype(np.float32),
                                 np.asarray(y, np.float64).astype(np.float32))

    def cbOdom(self, msg, delta_time=1.0):
        """msg: Odometry. delta_time replaces the wall-clock difference of ekf.cpp:80 (capped at 1).
        Returns the Landmarks message (x[], y[]) the reference publishes, or None while gated."""
        lib = capi.lib()
        yaw = lib.aslam_quat2euler(msg.qw, msg.qx, msg.qy, msg.qz)
        dt = np.float32(min(float(delta_time), 1.0))
        try:
            self._f.step([msg.x], [msg.y], [yaw], [msg.vx], [msg.az], [dt])
        except capi.AslamError as e:
            if e.status == capi.ASLAM_ERR_NOT_READY:  # `if (init_z) return;` ekf.cpp:76
                return None
            raise
        lx, ly, c = self._f.get_landmarks()
        n = int(c[0])
        return lx[0, :n].copy(), ly[0, :n].copy()

    # state accessors (param.X / param.P / N of the reference)
    def state(self):
        s = self._f.get_state()
        n = int(s["N"][0])
        return {"N": n, "X": s["X"][0, :n].copy(), "P": s["P"][0, :n, :n].copy(),
                "status": int(s["status"][0]), "assoc": s["assoc"].copy()}

    @property
    def batched(self):
        return self._f


class EKFSlam(_SingleFilter):
    """aslam::EKFSlam (src/ekf/ekf.h:71-132)."""
    KIND = "ekf"


class UKFSlam(_SingleFilter):
    """aslam::UKFSlam (src/ukf/ukf.h:86-144)."""
    KIND = "ukf"

    def llt_fail(self):
        return int(self._f.get_llt_fail()[0])


class LargeFilter:
    """One large-map filter (aslam_large handle): n = 3 + 2 L, P is n x n FP64 in HBM.

    EKF: rows [row_begin, row_end) of P live on this handle's GPU (a full handle has 0, n).
    """

    def __init__(self, kind, n, device=0, row_begin=0, row_end=None, config=None, **cfg_overrides):
        self._h = C.c_void_p()
        self._lib = capi.lib()
        self.config = config if config is not None else capi.default_config(**cfg_overrides)
        self.kind, self.n, self.device = kind, int(n), device
        self.row_begin = int(row_begin)
        self.row_end = int(n if row_end is None else row_end)
        if kind == "ekf":
            capi.check(self._lib.aslam_large_ekf_create(C.byref(self.config), self.n, self.row_begin, self.row_end,
                                                        device, C.byref(self._h)))
        else:
            capi.check(self._lib.aslam_large_ukf_create(C.byref(self.config), self.n, device, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.aslam_large_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self._lib.aslam_large_stream(self._h)

    def sync(self):
        capi.check(self._lib.aslam_large_sync(self._h))

    def load_state(self, X, P, Z=None):
        """X [n]; P [n, n] indexed P[i, j] (any layout: it is sent column-major); Z [n] (UKF)."""
        X = capi.as_np(X, np.float64)
        Pc = np.ascontiguousarray(np.asarray(P, np.float64).T)  # Pc[j, i] = P[i, j]: column-major
        Zc = None if Z is None else capi.as_np(Z, np.float64)
        capi.check(self._lib.aslam_large_load_state(self._h, capi.ptr(X), capi.ptr(Zc), capi.ptr(Pc), self.n))

    def load_state_colmajor(self, X, Pc, Z=None):
        """Pc: C array with Pc[j, i] = P[i, j] (no transpose copy)."""
        X = capi.as_np(X, np.float64)
        assert Pc.flags["C_CONTIGUOUS"] and Pc.dtype == np.float64
        Zc = None if Z is None else capi.as_np(Z, np.float64)
        capi.check(self._lib.aslam_large_load_state(self._h, capi.ptr(X), capi.ptr(Zc), capi.ptr(Pc), self.n))

    def get_state(self, want_P=True):
        """Returns X [n] and this handle's rows of P as an array [rows, n] (P[i - row_begin, j])."""
        X = np.zeros(self.n, np.float64)
        rows = self.row_end - self.row_begin
        Pc = np.zeros((self.n, rows), np.float64) if want_P else None
        capi.check(self._lib.aslam_large_get_state(self._h, capi.ptr(X), capi.ptr(Pc), rows))
        return X, (Pc.T if want_P else None)

    def step(self, x, y, yaw, vx, az, dt, rng, brg, want_assoc=True):
        rng = capi.as_np(rng, np.float32)
        brg = capi.as_np(brg, np.float32)
        assoc = np.zeros(max(rng.size, 1), np.int32) if want_assoc else None
        capi.check(self._lib.aslam_large_step(self._h, x, y, np.float32(yaw), vx, az, np.float32(dt), capi.ptr(rng),
                                              capi.ptr(brg), rng.size, capi.ptr(assoc)))
        return None if assoc is None else assoc[:rng.size]

    # sharded EKF: phase1 -> (caller all-reduces exchange buffer) -> phase2
    def step_phase1(self, x, y, yaw, vx, az, dt, rng, brg):
        rng = capi.as_np(rng, np.float32)
        brg = capi.as_np(brg, np.float32)
        q = C.c_uint32()
        capi.check(self._lib.aslam_large_step_phase1(self._h, x, y, np.float32(yaw), vx, az, np.float32(dt),
                                                     capi.ptr(rng), capi.ptr(brg), rng.size, C.byref(q)))
        self._m = rng.size
        return int(q.value)

    def exchange_buffer(self):
        """(device pointer, count of doubles) of the buffer the ranks must sum."""
        p = C.c_void_p()
        cnt = C.c_size_t()
        capi.check(self._lib.aslam_large_exchange_buffer(self._h, C.byref(p), C.byref(cnt)))
        return int(p.value), int(cnt.value)

    def step_phase2(self, want_assoc=True):
        assoc = np.zeros(max(self._m, 1), np.int32) if want_assoc else None
        capi.check(self._lib.aslam_large_step_phase2(self._h, capi.ptr(assoc)))
        return None if assoc is None else assoc[:self._m]

    # sharded EKF, fused exchange: peer stores over NVLink + arrival counters, no collective call
    def peer_export(self):
        """(64-byte CUDA IPC handle, raw device pointer) of this handle's exchange block."""
        hd = (C.c_ubyte * 64)()
        blk = C.c_void_p()
        capi.check(self._lib.aslam_large_peer_export(self._h, C.cast(hd, C.c_void_p), C.byref(blk)))
        return bytes(hd), int(blk.value)

    def peer_attach(self, world, rank, ipc_handles=None, blocks=None):
        """ipc_handles: list of `world` 64-byte handles (ranks in other processes) or blocks: list of `world`
        raw device pointers (ranks in this process). All ranks attach, then barrier, before the first step."""
        hbuf, barr = None, None
        if ipc_handles is not None:
            assert len(ipc_handles) == world and all(len(b) == 64 for b in ipc_handles)
            hbuf = C.create_string_buffer(b"".join(ipc_handles), 64 * world)
        if blocks is not None:
            assert len(blocks) == world
            barr = (C.c_void_p * world)(*blocks)
        capi.check(self._lib.aslam_large_peer_attach(self._h, world, rank, C.cast(hbuf, C.c_void_p) if hbuf else None,
                                                     barr))

    def step_fused(self, x, y, yaw, vx, az, dt, rng, brg, want_assoc=True):
        rng = capi.as_np(rng, np.float32)
        brg = capi.as_np(brg, np.float32)
        self._m = rng.size
        assoc = np.zeros(max(self._m, 1), np.int32) if want_assoc else None
        capi.check(self._lib.aslam_large_step_fused(self._h, x, y, np.float32(yaw), vx, az, np.float32(dt),
                                                    capi.ptr(rng), capi.ptr(brg), rng.size, capi.ptr(assoc)))
        return None if assoc is None else assoc[:self._m]

    def set_precision(self, mode):
        """UKF: precision of the three sigma-point products: 0 FP64 (default), 1 TF32, 3 = 3xTF32 (tcgen05 + TMEM)."""
        capi.check(self._lib.aslam_large_ukf_set_precision(self._h, int(mode)))

    def products_ms(self):
        """UKF: CUDA-event time of the three sigma-point products of the last step (TF32 modes: split included)."""
        ms = C.c_float()
        capi.check(self._lib.aslam_large_ukf_products_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def set_timing(self, on=True):
        """EKF: record CUDA events around the streaming update of every later step (off by default: the event keeps
        the update from being launched early behind the solve kernel)."""
        capi.check(self._lib.aslam_large_set_timing(self._h, 1 if on else 0))

    def last_kernel_ms(self):
        ms = C.c_float()
        capi.check(self._lib.aslam_large_last_kernel_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def status(self):
        st = C.c_uint32()
        fk = C.c_int32()
        rc = self._lib.aslam_large_status(self._h, C.byref(st), C.byref(fk))
        if rc != capi.ASLAM_ERR_TIMEOUT:  # a peer timeout is reported through the status bit as well
            capi.check(rc)
        return int(st.value), int(fk.value)

    @property
    def launch_count(self):
        return int(self._lib.aslam_large_launch_count(self._h))


class LandMarks:
    """Host mirror of the reference's sensor_landmark node (aslam::LandMarks, sensor_landmark.h:49-89): `callback`
    takes LaserScans (360 ranges each, +inf = no return) and returns what the node publishes on
    out/landmarks/sensor -- per scan the (range, bearing) lists of the fitted cylinders -- computed by the CUDA
    kernel of csrc/laser_front.cu through the C ABI."""

    MAX_LANDMARKS = 32

    def __init__(self, max_scans=1, device=0):
        self._lib = capi.lib()
        self._h = C.c_void_p()
        self.max_scans = int(max_scans)
        capi.check(self._lib.aslam_laser_create(None, self.max_scans, device, C.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.aslam_laser_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    @property
    def stream(self):
        return self._lib.aslam_laser_stream(self._h)

    @property
    def launch_count(self):
        return int(self._lib.aslam_laser_launch_count(self._h))

    def sync(self):
        capi.check(self._lib.aslam_laser_sync(self._h))

    def callback(self, ranges):
        """ranges: [nscan, 360] float32 (or [360]). Returns (range [nscan, 32], bearing [nscan, 32], counts [nscan]);
        counts = -1 where the reference's assert(theta < 359) fires."""
        ranges = capi.as_np(ranges, np.float32).reshape(-1, 360)
        ns = ranges.shape[0]
        rg = np.zeros((ns, self.MAX_LANDMARKS), np.float32)
        bg = np.zeros((ns, self.MAX_LANDMARKS), np.float32)
        cnt = np.zeros(ns, np.int32)
        capi.check(self._lib.aslam_laser_extract(self._h, capi.ptr(ranges), ns, capi.ptr(rg), capi.ptr(bg), capi.ptr(cnt)))
        return rg, bg, cnt

    def callback_into(self, ranges, range_out, bearing_out, counts):
        """`callback` into caller-owned host arrays (e.g. views of pinned memory: the copies then overlap the kernels):
        ranges [nscan, 360] float32; range_out / bearing_out [nscan, 32] float32; counts [nscan] int32, all C-contiguous."""
        ns = ranges.shape[0]
        for a, dt, shape in ((ranges, np.float32, (ns, 360)), (range_out, np.float32, (ns, self.MAX_LANDMARKS)),
                             (bearing_out, np.float32, (ns, self.MAX_LANDMARKS)), (counts, np.int32, (ns,))):
            if a.dtype != dt or a.shape != shape or not a.flags["C_CONTIGUOUS"]:
                raise ValueError("callback_into: wrong dtype / shape / layout")
        capi.check(self._lib.aslam_laser_extract(self._h, capi.ptr(ranges), ns, capi.ptr(range_out), capi.ptr(bearing_out),
                                                 capi.ptr(counts)))

    def synthesize(self, lmx, lmy, poses, noise=None, sigma=0.01, radius=0.25):
        """Synthetic LDS scans on the GPU (the Gazebo stand-in): poses [nscan, 3] (x, y, yaw), noise [nscan, 360]
        standard-normal float32 or None -> ranges [nscan, 360] float32 (+inf = no return)."""
        lmx = capi.as_np(lmx, np.float64)
        lmy = capi.as_np(lmy, np.float64)
        poses = capi.as_np(poses, np.float64).reshape(-1, 3)
        ns = poses.shape[0]
        if noise is not None:
            noise = capi.as_np(noise, np.float32).reshape(ns, 360)
        out = np.zeros((ns, 360), np.float32)
        capi.check(self._lib.aslam_laser_synthesize(self._h, capi.ptr(lmx), capi.ptr(lmy), lmx.size, capi.ptr(poses),
                                                    capi.ptr(noise), ns, float(sigma), float(radius), capi.ptr(out)))
        return out

    def callback_device(self, d_ranges, nscan, d_range_out, d_bearing_out, d_counts):
        """Device pointers (ints); asynchronous on the handle's stream."""
        capi.check(self._lib.aslam_laser_extract_device(self._h, d_ranges, nscan, d_range_out, d_bearing_out, d_counts))
