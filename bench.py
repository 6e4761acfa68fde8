#!/usr/bin/env python
"""bench.py -- EKF-SLAM filter updates/s of the B200 path (BASELINE.json metric), one JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  (N > 1: python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...)

Workload (config.workload = "cfg3"): 65,536 independent EKF-SLAM filters x 16 landmarks PER GPU
(weak scaling; the filters are independent, so the batch shards with no collective), synthetic
replay of SURVEY.md section 8d, per-filter seeds, every filter already at N = 35 when the timed region
starts. One bench "step" = one launch that replays FUSE consecutive messages for every filter of
the batch (the covariance stays on-chip between them); updates = filters x FUSE per step.

  value     updates/s, messages resident in HBM (aslam_filter_run_device)
  e2e       same through the host-buffer entry point (aslam_filter_run_host from pinned memory +
            a device->host read of param.X every step)
  roofline  the HBM-bound kernel north_star sets a roofline target for: the large-map (cfg4,
            n = 8195) streaming covariance update, timed live with CUDA events on its own stream in
            this same run (second leg, rank 0 / sharded over the ranks when N > 1)
  roofline_batched   the cfg3 kernel against the measured FP64-pipe peak and against HBM
  cpu_baseline       the CPU oracle (port of the reference math) on the host cores, bounded sample

The reference arm (--impl reference) times that same oracle with all host threads on the same
config; the reference itself needs ROS + Eigen and cannot be built here.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU = 65536
L_CFG3 = 16
FUSE = 8               # messages per launch
WARM_MSGS = 24         # replayed before timing: 15 stationary + promotion on the 10th + settle
FLOP_PER_UPDATE = 6 * 35 ** 3          # SURVEY.md section 8d convention (6 n^3), fixed
BYTES_PER_UPDATE = 20888               # X, Z, P read + written, inputs


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=B_PER_GPU, help="filters per GPU")
    ap.add_argument("--no-large", action="store_true", help="skip the cfg4 large-map leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU-baseline leg")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for k, nme in enumerate(names):
                if r[5 + k].lower().startswith("active"):
                    reasons.add(nme)
        sm_sorted = sorted(sm)
        # samples taken while the GPU was busy are the upper part of the distribution
        load = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        return {"sm_mhz": statistics.median(load) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------ CPU legs
def host_threads():
    """Host threads the CPU legs use: every core this process may run on. Passed explicitly because
    torch.distributed.run exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_filters_run(kind, rp, B, T, L, nthr, cap64):
    """B filters x T messages of replay `rp` (from reset) on `nthr` host threads. Uses oracle/_ref -- the reference's
    own ekf.cpp / ukf.cpp compiled unmodified (cap64: with MAX_LANDMARK_COUNT lifted to 64 at build time) -- when it
    was built ("reference"), else the restated oracle ("port"). Returns (seconds, N per filter, kind)."""
    from oracle import ref as R
    sl = (slice(0, T), slice(0, B))
    if R.available(kind, cap64):
        sec, X, P, N = R.run_batch(kind, B, T, L, rp.x[sl], rp.y[sl], rp.qw[sl], rp.qz[sl], rp.vx[sl], rp.az[sl],
                                   rp.range[sl], rp.bearing[sl], nthr, 36, want_P=False, cap64=cap64)
        return sec, N, "reference"
    from oracle import oracle as O
    O.build()
    cfg = O.default_config(max_state_dim=64 if cap64 else 30)
    sec, X, P, N = O.run_batch(kind, cfg, B, T, L, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl],
                               rp.range[sl], rp.bearing[sl], nthr, 36, want_P=False)
    return sec, N, "port"


CPU_KIND_NOTE = {
    "reference": "oracle/_ref: the reference's own ekf.cpp/ukf.cpp + headers compiled unmodified (g++ -O2) against "
                 "Eigen/ROS stand-ins; MAX_LANDMARK_COUNT 30 -> 64 at build time for n = 35",
    "port": "oracle/: restated port of the reference math (oracle/_ref not built)",
}


def cpu_leg(sample_filters, msgs, threads=0):
    """The reference's CPU filter on `sample_filters` cfg3 filters x `msgs` messages from reset, all host threads,
    plus the same on ONE thread over a smaller sample. Returns the cpu_baseline object."""
    import awesomeslam_b200 as A
    lmx, lmy = A.map_cfg3()
    rp = A.Replay(lmx, lmy, sample_filters, msgs, seed0=A.replay.SEED_CFG3)
    nthr = threads or host_threads()
    sec, N, kind = cpu_filters_run("ekf", rp, sample_filters, msgs, L_CFG3, nthr, True)
    assert np.all(N == 35), "CPU filters did not reach N = 35"
    one_b = max(16, sample_filters // (2 * nthr))
    sec1, N1, _ = cpu_filters_run("ekf", rp, one_b, msgs, L_CFG3, 1, True)
    return {"value": sample_filters * msgs / sec, "unit": "updates/s", "cores": nthr, "kind": kind,
            "sample": "%d of the 65,536 cfg3 filters x %d messages from reset, OpenMP over filters, %.1f s"
                      % (sample_filters, msgs, sec),
            "one_core": {"value": one_b * msgs / sec1, "unit": "updates/s", "cores": 1,
                         "sample": "%d filters x %d messages, %.1f s" % (one_b, msgs, sec1)},
            "what": CPU_KIND_NOTE[kind]}


def reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path -- oracle/_ref, i.e. ekf.cpp and its
    headers compiled unmodified against the Eigen / ROS stand-ins (the restated oracle port only where that was not
    built) -- with all host threads, on the cfg3 workload. One step = FUSE messages for a bounded sample of the
    batch. Nothing of the product's CUDA library is loaded (the replay generator is lib/libaslam_host.so)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import awesomeslam_b200 as A
    nthr = host_threads()
    sample = 256 if nthr < 16 else 2048
    lmx, lmy = A.map_cfg3()
    total_msgs = WARM_MSGS + (args.warmup + args.steps) * FUSE
    rp = A.Replay(lmx, lmy, sample, total_msgs, seed0=A.replay.SEED_CFG3)
    # every filter starts from reset (the reference has no load_state): the whole replay is timed and attributed to
    # the steps in proportion to their messages. The warm-up messages at N = 3 are cheaper than the N = 35 ones the
    # GPU arm times, so this slightly FAVOURS the CPU; filters are independent, so the sample extrapolates exactly.
    sec, N, kind = cpu_filters_run("ekf", rp, sample, total_msgs, L_CFG3, nthr, True)
    assert np.all(N == 35)
    ups = sample * total_msgs / sec
    ms_per_step = 1e3 * sec * (FUSE / total_msgs)
    line = {
        "impl": "reference", "metric": "ekf_slam_filter_updates_per_s", "value": ups, "unit": "updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "cfg3: 65,536 EKF-SLAM filters x 16 landmarks per GPU (n=35)",
                   "sample": "%d filters x %d messages" % (sample, total_msgs), "fuse": FUSE},
        "cpu_baseline": {"value": ups, "unit": "updates/s", "cores": nthr, "kind": kind,
                         "sample": "%d of 65,536 filters x %d messages from reset, OpenMP over filters"
                                   % (sample, total_msgs), "what": CPU_KIND_NOTE[kind]},
        "e2e": {"value": ups, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# ------------------------------------------------------------------------------ GPU legs
class _DevArray:
    """CUDA-array-interface view of library-owned device memory (for torch.as_tensor)."""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 2}


def batched_leg(args, torch, dist, rank, world, dev):
    import awesomeslam_b200 as A
    B = args.batch
    K, W = args.steps, args.warmup
    lmx, lmy = A.map_cfg3()
    total_msgs = WARM_MSGS + (W + K) * FUSE
    # per-filter seeds 0x5EED0000 + global filter id (rank-major)
    rp = A.Replay(lmx, lmy, B, total_msgs, seed0=A.replay.SEED_CFG3, first_filter=rank * B)
    host = [rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing]
    dev_t = [torch.from_numpy(a).to(dev) for a in host]
    pinned = [torch.from_numpy(a).pin_memory() for a in host]

    def dptrs(msg0):
        out = []
        for t in dev_t:
            per = t[0].numel() * t.element_size()
            out.append(t.data_ptr() + msg0 * per)
        return out

    def run_resident(f, msg0, nmsg):
        p = dptrs(msg0)
        f.run_device(nmsg, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], L_CFG3)

    def run_hostbuf(f, msg0, nmsg):
        sl = slice(msg0, msg0 + nmsg)
        a = [t[sl].numpy() for t in pinned]
        f.run_host(nmsg, a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], L_CFG3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    results = {}
    for mode in ("resident", "e2e"):
        f = A.BatchedFilter("ekf", B, dev.index, max_state_dim=36)
        stream = torch.cuda.ExternalStream(f.stream, device=dev)
        xouts = [torch.empty((B, f.ld), dtype=torch.float64).pin_memory() for _ in range(2)]
        xout = xouts[0]
        run_resident(f, 0, WARM_MSGS)  # bring every filter to N = 35 (not timed)
        f.sync()
        st = f.get_state(want_P=False, want_assoc=False)
        assert np.all(st["N"] == 35) and not np.any(st["status"]), "warm-up did not map all 16 landmarks"
        runner = run_resident if mode == "resident" else run_hostbuf
        msg = WARM_MSGS
        for _ in range(W):
            runner(f, msg, FUSE)
            if mode == "e2e":
                f.read_state_into(xout.numpy())
            msg += FUSE
        l0 = f.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(dev.index)
        if mode == "resident" and rank == 0:
            sampler.start()
        barrier()
        # e2e: the pipelined public API. Every step copies its 8 messages host->device (second stream, overlapping
        # the previous step's kernel) and reads param.X back to pinned host memory; the host consumes result k-1
        # while step k runs, and waits for the last read before the closing event is recorded.
        chk, prev = 0.0, None
        with torch.cuda.stream(stream):
            e0.record()
            for k in range(K):
                runner(f, msg, FUSE)
                if mode == "e2e":
                    t = f.read_state_async(xouts[k & 1].numpy())
                    if prev is not None:
                        f.read_wait(prev)
                        chk += float(xouts[(k - 1) & 1][0, 0])
                    prev = t
                msg += FUSE
            if mode == "e2e":
                f.read_wait(prev)
                chk += float(xouts[(K - 1) & 1][0, 0])
            e1.record()
        assert np.isfinite(chk)
        f.sync()
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if (mode == "resident" and rank == 0) else None
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        st = f.get_state(want_P=False, want_assoc=False)
        assert np.all(st["N"] == 35) and not np.any(st["status"]) and np.all(np.isfinite(st["X"]))
        results[mode] = {"ms": ms, "launches": f.launch_count - l0, "clocks": clocks, "ld": f.ld}
        if mode == "resident" and rank == 0 and not args.no_cpu:
            results["parity"] = cfg3_parity(f, rp, B, total_msgs)
        f.close()
    per_msg = rp.nbytes_per_step()
    results["h2d_bytes_per_step"] = per_msg * FUSE
    results["d2h_bytes_per_step"] = B * results["e2e"]["ld"] * 8
    return results


def cfg3_parity(f, rp, B, total_msgs):
    """Driver-visible parity of the TIMED batch: after the timed region the state of 64 filters
    spread over the batch -- first / last slot of the first / last CTA (7 filters per CTA) and 58 more -- is compared
    with the CPU filter (oracle/_ref when built: the compiled reference) replaying the same messages from reset."""
    from oracle import ref as R
    ids = sorted(set([0, 6, 7, 13, B - 1, B - 2, (B - 1) // 7 * 7, (B - 1) // 7 * 7 - 1]
                     + [int(i) for i in np.linspace(14, B - 3, 56)]))
    ids = np.array(ids, dtype=np.int64)
    st = f.get_state(want_P=True, want_assoc=False)
    sl = (slice(0, total_msgs), ids)
    c = lambda a: np.ascontiguousarray(a[sl])
    if R.available("ekf", True):
        sec, X, P, N = R.run_batch("ekf", ids.size, total_msgs, L_CFG3, c(rp.x), c(rp.y), c(rp.qw), c(rp.qz), c(rp.vx),
                                   c(rp.az), c(rp.range), c(rp.bearing), host_threads(), 36, want_P=True, cap64=True)
        kind = "reference"
    else:
        from oracle import oracle as O
        O.build()
        sec, X, P, N = O.run_batch("ekf", O.default_config(max_state_dim=64), ids.size, total_msgs, L_CFG3, c(rp.x),
                                   c(rp.y), c(rp.yaw), c(rp.vx), c(rp.az), c(rp.dt), c(rp.range), c(rp.bearing),
                                   host_threads(), 36, want_P=True)
        kind = "port"
    ex = [float(np.max(np.abs(st["X"][i] - X[k])) / np.max(np.abs(X[k]))) for k, i in enumerate(ids)]
    ep = [float(np.max(np.abs(st["P"][i] - P[k])) / np.max(np.abs(P[k]))) for k, i in enumerate(ids)]
    assert np.array_equal(st["N"][ids], N)
    worst = max(max(ex), max(ep))
    assert worst <= 1e-9, "timed cfg3 batch differs from the CPU filter: %.3e" % worst
    return {"filters_compared": int(ids.size), "filter_ids": "0, 6, 7, 13, ..., %d, %d (first/last slots of the first/"
            "last CTA + 56 spread over the batch)" % (B - 2, B - 1), "messages": int(total_msgs), "against": kind,
            "worst_relerr_X": max(ex), "worst_relerr_P": max(ep),
            "filters_beyond_1e-9": int(sum(1 for a, b in zip(ex, ep) if max(a, b) > 1e-9)), "tolerance": 1e-9}


def large_leg(args, torch, dist, rank, world, dev, hbm_peak, peak_src):
    """cfg4: n = 8195, 8 observed landmarks per step (q = 19). Rows of P sharded over the ranks;
    the exchange buffer (2 q n doubles) is summed with one NCCL all-reduce per step when world > 1."""
    import awesomeslam_b200 as A
    K, W = args.steps, args.warmup
    tx, ty, X0, G, d = A.large_map_state(64)
    n = X0.size
    P0 = A.dense_P(G, d)
    rb = 0 if rank == 0 else max(3, (n * rank) // world)
    re = n if rank == world - 1 else max(3, (n * (rank + 1)) // world)
    f = A.LargeFilter("ekf", n, dev.index, row_begin=rb, row_end=re)
    f.load_state_colmajor(X0, P0)
    del P0
    rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
    stream = torch.cuda.ExternalStream(f.stream, device=dev)
    exch = None
    ODOM = (0.0, 0.0, 0.0, 0.05, 0.1, 1.0)

    def step_single():
        f.step(*ODOM, rng, brg, want_assoc=False)

    def step_nccl():
        nonlocal exch
        f.step_phase1(*ODOM, rng, brg)
        if exch is None:
            p, c = f.exchange_buffer()
            exch = torch.as_tensor(_DevArray(p, c), device=dev)
        dist.all_reduce(exch)
        f.step_phase2(want_assoc=False)

    def step_fused():
        f.step_fused(*ODOM, rng, brg, want_assoc=False)

    def timed(step):
        with torch.cuda.stream(stream):
            f.set_timing(False)
            for _ in range(max(W, 3)):
                step()
            f.sync()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = f.launch_count
            KL = max(K, 400)  # steps of 0.05-0.2 ms: enough of them that neither the fill of the step pipeline nor the
            # skew between the ranks leaving the barrier (~1 ms, inside the slowest rank's window) shows in the figure
            e0.record()
            th0 = time.perf_counter()
            for _ in range(KL):
                step()
            host_ms = (time.perf_counter() - th0) * 1e3 / KL  # host time to ENQUEUE a step (no synchronisation inside)
            e1.record()
            f.sync()
            torch.cuda.synchronize(dev)
            ms = e0.elapsed_time(e1) / KL
            launches = (f.launch_count - l0) * K / KL
            # the streaming kernel alone: CUDA events around it on the handle's stream (a separate pass: the event
            # between the solve kernel and the update disables the update's early launch, so those steps are slower)
            f.set_timing(True)
            kms = []
            for _ in range(K):
                step()
                kms.append(f.last_kernel_ms())
            f.sync()
            f.set_timing(False)
        if world > 1:
            dist.barrier()
        return ms, launches, kms, host_ms

    exchange, nccl_ms, parity = "none (one GPU)", None, None
    if world == 1:
        ms, launches, kms, host_ms = timed(step_single)
    else:
        # comparison point: phase1 -> NCCL all-reduce of the exchange buffer -> phase2
        nccl_ms, launches, kms, host_ms = timed(step_nccl)
        ms, exchange = nccl_ms, "NCCL all-reduce of the [2q][n] exchange buffer"
        try:  # the fused exchange: CUDA IPC peer mappings, stores + arrival counters inside the solve kernel
            hd, _ = f.peer_export()
            handles = [None] * world
            dist.all_gather_object(handles, hd)
            f.peer_attach(world, rank, ipc_handles=handles)
            dist.barrier()
            ms, launches, kms, host_ms = timed(step_fused)
            if os.environ.get("ASLAM_LARGE_TRACE"):  # diagnostic: %globaltimer phases of the last steps, to stderr
                tr = np.zeros((64, 16), np.uint64)
                A.capi.check(f._lib.aslam_large_trace(f._h, A.capi.ptr(tr)))
                tr = tr.astype(np.int64)
                ep = int(np.argmax(tr[:, 2] * (tr[:, 2] < (1 << 62))))
                rows_ = []
                for kk in range(ep - 32, ep - 18):  # (the last K steps are the serialised kernel-timing pass)
                    a_, b_ = tr[kk % 64], tr[(kk + 1) % 64]
                    rows_.append(dict(solve=(a_[3] - a_[2]) / 1e3, gathered=(a_[8] - a_[2]) / 1e3, psub=(a_[9] - a_[8]) / 1e3,
                                      upd=(a_[5] - a_[4]) / 1e3, solve_out_to_upd_go=(a_[4] - a_[3]) / 1e3,
                                      upd_out_to_next_solve_in=(b_[2] - a_[5]) / 1e3, period=(b_[2] - a_[2]) / 1e3))
                per_ = [int((tr[(kk + 1) % 64][2] - tr[kk % 64][2]) / 1e3) for kk in range(ep - 52, ep)]
                sys.stderr.write("rank %d periods (us) of the last 52 steps: %s\n" % (rank, per_))
                sys.stderr.write("rank %d trace: %s\n" % (rank, "  ".join(
                    "%s %.1f" % (k_, float(np.median([r_[k_] for r_ in rows_]))) for k_ in rows_[0])))
            assert f.status()[0] == 0, "fused exchange timed out"
            exchange = ("fused: peer stores over NVLink + arrival counters inside the solve kernel "
                        "(no collective call, 3 launches per step)")
            parity = large_parity(f, step_nccl, step_fused, torch, dist, rank, world, dev, rb, re, X0, G, d,
                                  rng, brg, ODOM, stream)
        except AssertionError:
            raise
        except Exception as e:  # noqa: BLE001  (reported, and the NCCL numbers stand)
            sys.stderr.write("fused exchange unavailable: %r\n" % (e,))
    t = torch.tensor([ms, float(np.mean(kms))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tn = torch.tensor([nccl_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tn, op=dist.ReduceOp.MAX)
        nccl_ms = float(tn.item())
    ms, kernel_ms = float(t[0].item()), float(t[1].item())
    X, _ = f.get_state(want_P=False)
    assert np.all(np.isfinite(X))
    f.close()
    rows = re - rb
    bytes_launch = 2 * 8 * rows * n  # this rank's rows of P, read + written once
    achieved = bytes_launch / (kernel_ms * 1e-3) / 1e9
    return {
        "workload": "cfg4: large-map EKF n=8195 (4096 landmarks), q=19 rows of H per step, FP64 P%s"
                    % ("" if world == 1 else ", rows sharded over %d GPUs, one exchange per step" % world),
        "exchange": exchange, "ms_per_step_nccl_exchange": nccl_ms, "parity": parity,
        "steps_per_s": 1e3 / ms, "ms_per_step": ms, "launches_per_step": launches / K, "steps_timed": max(K, 400),
        "host_enqueue_ms_per_step": host_ms,
        "roofline": {"bound": "hbm", "kernel": "large_update_stream_kernel<7,2,20,12,4> (persistent, 1 CTA/SM)", "achieved": achieved,
                     "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                     "traffic": _ncu_traffic() if world == 1 else None,
                     "algorithmic_bytes_per_launch": bytes_launch, "kernel_ms": kernel_ms, "peak_source": peak_src},
    }


def large_parity(f, step_nccl, step_fused, torch, dist, rank, world, dev, rb, re, X0, G, d, rng, brg, ODOM, stream,
                 steps=3):
    """Driver-visible parity of the multi-process exchange (bit-identical by design, asserted here on hardware):
    from the same loaded state, `steps` steps through the NCCL all-reduce form and through the fused peer-store form
    must leave EQUAL X and EQUAL local rows of P on every rank; and rank 0 runs the same steps on an UNSHARDED handle
    of its own GPU and compares X and, row block by row block (gathered over the ranks), the whole P."""
    import awesomeslam_b200 as A
    n = X0.size
    P0 = A.dense_P(G, d)
    out = {}
    res = {}
    for name, step in (("nccl", step_nccl), ("fused", step_fused)):
        f.load_state_colmajor(X0, P0)
        if world > 1:
            dist.barrier()
        with torch.cuda.stream(stream):  # the all-reduce of the NCCL form must be ordered on the handle's stream
            for _ in range(steps):
                step()
        f.sync()
        torch.cuda.synchronize(dev)
        res[name] = f.get_state()
        if world > 1:
            dist.barrier()
    eq = bool(np.array_equal(res["nccl"][0], res["fused"][0]) and np.array_equal(res["nccl"][1], res["fused"][1]))
    flag = torch.tensor([1 if eq else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out["fused_eq_nccl"] = bool(flag.item())
    # against the unsharded handle on rank 0 (1 GB more on that GPU, freed at once)
    digest = torch.zeros(2, dtype=torch.float64, device=dev)
    Pl = res["fused"][1]  # [rows, n]
    if rank == 0:
        u = A.LargeFilter("ekf", n, dev.index)
        u.load_state_colmajor(X0, P0)
        for _ in range(steps):
            u.step(*ODOM, rng, brg, want_assoc=False)
        Xu, Pu = u.get_state()
        u.close()
        bounds = [None] * world
    else:
        Xu = Pu = None
        bounds = None
    del P0
    # every rank sends its rows to rank 0 in turn (gather_object of 67 MB blocks at N = 8: bounded, once per run)
    ok = True
    for src in range(world):
        if src == 0:
            blk = Pl if rank == 0 else None
            lo, hi = (rb, re) if rank == 0 else (0, 0)
        else:
            if rank == src:
                t = torch.from_numpy(np.ascontiguousarray(Pl)).to(dev)
                meta = torch.tensor([rb, re], dtype=torch.int64, device=dev)
                dist.send(meta, 0)
                dist.send(t, 0)
            if rank == 0:
                meta = torch.zeros(2, dtype=torch.int64, device=dev)
                dist.recv(meta, src)
                lo, hi = int(meta[0].item()), int(meta[1].item())
                t = torch.empty((hi - lo, n), dtype=torch.float64, device=dev)
                dist.recv(t, src)
                blk = t.cpu().numpy()
        if rank == 0:
            ok = ok and bool(np.array_equal(blk, Pu[lo:hi]))
    if rank == 0:
        ok = ok and bool(np.array_equal(res["fused"][0], Xu))
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    dist.broadcast(flag, 0)
    out["eq_unsharded"] = bool(flag.item())
    out["steps"] = steps
    assert out["fused_eq_nccl"] and out["eq_unsharded"], "sharded large-map EKF differs from the unsharded one: %r" % out
    return out


def _ncu_traffic(kernel="large_update_stream_kernel"):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the cfg4 streaming update, from the NEWEST committed
    `ncu --set full` capture (profiles/rNN_traffic.json names the capture and the kernel it was taken from). None when
    no file is there or when the newest capture is of another kernel (a stale number is worse than none)."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r[0-9][0-9]_traffic.json")))
    if not files:
        return None
    try:
        with open(files[-1]) as fh:
            t = json.load(fh)
        if kernel not in t.get("kernel", ""):
            return None
        return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
    except (OSError, ValueError, KeyError):
        return None


def _ncu_batched_pipe():
    """FP64-pipe utilisation of the cfg3 batched kernel from the newest committed ncu capture
    (profiles/rNN_batched_pipe.json, written by tools/collect_profiles.py); None when absent."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r[0-9][0-9]_batched_pipe.json")))
    if not files:
        return None
    try:
        with open(files[-1]) as fh:
            return json.load(fh).get("pipe_fp64_active_pct")
    except (OSError, ValueError):
        return None


def ukf_batched_leg(args, torch, dev, B=16384, shape="cfg2"):
    """Batched UKF-SLAM (K2). shape "cfg2": the reference's UKFSlam at its own size -- the landmarks.yaml map, n = 19,
    43 sigma points, max_state_dim 30 (the reference's cap). shape "cfg3": 16 landmarks, n = 35, n_aug = 37, 75 sigma
    points, max_state_dim 36. B filters with their own noise seeds, messages resident in HBM, FUSE
    messages per launch. The free-running reference UKF turns indefinite (w0 < 0) and the kernel reproduces
    that, LLT failures included: the count of filters that reported one is printed, the time per update does not depend
    on it. CPU: the compiled reference (oracle/_ref) on a bounded sample."""
    import awesomeslam_b200 as A
    if shape == "cfg2":
        (lmx, lmy), msd, seed = A.map_cfg1(), 30, 0x0CF20000
    elif shape == "cfg2_cap20":  # the same filters under a cap of 20 (N = 19 still fits): smaller footprint, 6 per SM
        (lmx, lmy), msd, seed = A.map_cfg1(), 20, 0x0CF20000
    else:
        (lmx, lmy), msd, seed = A.map_cfg3(), 36, 0x0CF30000
    L = len(lmx)
    K, W = min(args.steps, 5 if shape == "cfg3" else 10), max(args.warmup, 3)
    total = WARM_MSGS + (W + K) * FUSE
    rp = A.Replay(lmx, lmy, B, max(total, 120), seed0=seed)
    dev_t = [torch.from_numpy(a).to(dev) for a in (rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing)]

    def run(f, msg0, nmsg):
        p = [t.data_ptr() + msg0 * t[0].numel() * t.element_size() for t in dev_t]
        f.run_device(nmsg, *p, L)

    f = A.BatchedFilter("ukf", B, dev.index, max_state_dim=msd)
    stream = torch.cuda.ExternalStream(f.stream, device=dev)
    run(f, 0, WARM_MSGS)
    f.sync()
    N0 = f.get_state(want_P=False, want_assoc=False)["N"]
    assert np.mean(N0 == 3 + 2 * L) > 0.99, "UKF warm-up did not map the landmarks"
    msg = WARM_MSGS
    for _ in range(W):
        run(f, msg, FUSE)
        msg += FUSE
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    with torch.cuda.stream(stream):
        e0.record()
        for _ in range(K):
            run(f, msg, FUSE)
            msg += FUSE
        e1.record()
    f.sync()
    ms = e0.elapsed_time(e1)
    st = f.get_state(want_P=False, want_assoc=False)
    failed = int(np.count_nonzero(st["status"] & A.capi.ST_LLT_FAILED))
    f.close()
    n = 3 + 2 * L
    ups = B * FUSE * K / (ms * 1e-3)
    out = {"workload": "batched UKF-SLAM: %d filters x %d landmarks (n=%d, %d sigma points, max_state_dim %d), "
                       "%d messages per launch" % (B, L, n, 2 * n + 5, msd, FUSE),
           "kernel": "ukf_batched_kernel<%d>" % msd,
           "updates_per_s": ups, "ms_per_launch": ms / K, "filters_with_llt_failure": failed,
           "fp64_tflops_20n3": ups * 20.0 * n ** 3 / 1e12}
    if not args.no_cpu and shape != "cfg2_cap20":
        nthr = host_threads()
        sample, msgs = ((1024 if nthr < 16 else 4096), 120) if shape == "cfg2" else ((256 if nthr < 16 else 1024), 80)
        sec, N, kind = cpu_filters_run("ukf", rp, sample, msgs, L, nthr, shape != "cfg2")
        out["cpu_baseline"] = {"value": sample * msgs / sec, "unit": "updates/s", "cores": nthr, "kind": kind,
                               "sample": "%d filters x %d messages from reset, OpenMP over filters, %.1f s" % (sample, msgs, sec)}
    return out


def large_cpu_leg(steps=10):
    """cfg4 on the host cores: the oracle's row-subset joint update (the same formula the kernels run), full size,
    OpenMP over the columns of P. Returns the cpu_baseline object of the large_map leg."""
    import time
    import awesomeslam_b200 as A
    from oracle import oracle as O
    O.build()
    tx, ty, X0, G, d = A.large_map_state(64)
    P = A.dense_P(G, d)
    X = X0.copy()
    rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
    cfg = O.default_config(max_state_dim=1 << 20)
    a00, a10 = 1.0, 0.0
    sec = []
    for it in range(steps + 1):
        t0 = time.perf_counter()
        q, assoc, a00, a10 = O.ekf_large_step(cfg, X, P, a00, a10, 0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, 0)
        sec.append(time.perf_counter() - t0)
    assert q == 19 and np.all(np.isfinite(X))
    t = float(np.mean(sec[1:]))
    out = {"value": 1.0 / t, "unit": "steps/s", "cores": O.num_threads(), "kind": "port",
           "sample": "%d full-size steps (n=8195, q=19) after one warm-up step, %.2f s each" % (steps, t)}
    # the reference's own form -- the verbatim DENSE EKFSlam::slam (ekf.cpp:293-311: N x N S, S.inverse(),
    # K = P H' S^-1, (I - K H) P) -- timed on one core at L = 121 and L = 256 landmarks and extrapolated by n^3 to
    # n = 8195 (labelled: an extrapolation; at full size the 537 MB matrices no longer fit any cache, so it favours the CPU)
    from oracle import np_restatement as R
    dense = []
    for side in (11, 16):
        tx, ty, X0, G, d = A.large_map_state(side)
        P0 = A.dense_P(G, d)
        rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
        o = O.Filter("ekf", max_state_dim=100000)
        o.load_state(X0, R.measurement(X0.size, X0), P0)
        o.set_observations(rng, brg)
        t0 = time.perf_counter()
        o.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0)
        dense.append((X0.size, time.perf_counter() - t0))
    n1, t1 = dense[-1]
    full = t1 * (8195.0 / n1) ** 3
    out["dense_reference_form"] = {
        "value": 1.0 / full, "unit": "steps/s", "cores": 1, "kind": "port", "extrapolated": True,
        "sample": "verbatim dense slam: n=%d %.2f s, n=%d %.2f s (exponent %.2f); x (8195/%d)^3 = %.0f s per full-size step"
                  % (dense[0][0], dense[0][1], n1, t1, np.log(t1 / dense[0][1]) / np.log(n1 / dense[0][0]), n1, full)}
    return out


def ukf_large_cpu_leg():
    """cfg5 on one host core: the oracle's verbatim UKFSlam::slam. A full-size step takes minutes (280 s measured in
    the build container), so the bounded sample is ONE step at n = 515 (256 landmarks, same recipe), scaled by n^3 --
    an extrapolation that favours the CPU (at full size the rank-1 sweeps no longer fit the caches)."""
    import time
    import awesomeslam_b200 as A
    from oracle import oracle as O, np_restatement as R
    O.build()
    side = 16
    tx, ty, X0, G, d = A.large_map_state(side, pitch=100.0 / side)
    P0 = A.dense_P(G, d)
    rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
    o = O.Filter("ukf", max_state_dim=100000)
    o.load_state(X0, R.measurement(X0.size, X0), P0)
    o.set_observations(rng, brg)
    t0 = time.perf_counter()
    o.step(0.01, -0.01, 0.002, 0.05, 0.1, 1.0)
    t = time.perf_counter() - t0
    assert o.llt_fail() == -1
    scale = (2051.0 / X0.size) ** 3
    return {"value": 1.0 / (t * scale), "unit": "steps/s", "cores": 1, "kind": "port",
            "sample": "one step at n=%d (%.2f s) x (2051/%d)^3 = %.0f s per full-size step (extrapolated)"
                      % (X0.size, t, X0.size, t * scale)}


def ukf_large_leg(args, torch, dev):
    """cfg5: large-map UKF, n = 2051 (1024 landmarks, 4107 sigma points), one message from the supplied positive
    definite state per step (a second consecutive message of the reference UKF already finds P indefinite, SURVEY 3.2:
    the state is re-loaded before every step, outside the timed region). Timed: the whole device chain of the step
    (Cholesky, sigma points, the three weighted-product GEMMs, solves, update) with CUDA events on the handle's stream."""
    import ctypes as C
    import awesomeslam_b200 as A
    side = 32
    tx, ty, X0, G, d = A.large_map_state(side, pitch=100.0 / side)
    P0 = A.dense_P(G, d)
    n = X0.size
    rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
    f = A.LargeFilter("ukf", n, dev.index)
    ms = []
    for it in range(max(args.warmup, 3) + min(args.steps, 10)):
        f.load_state(X0, P0)
        l0 = f.launch_count
        f.step(0.01, -0.01, 0.002, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
        t = f.last_kernel_ms()
        launches = f.launch_count - l0
        if it >= max(args.warmup, 3):
            ms.append(t)
    X, P = f.get_state()
    status, fail = f.status()
    assert np.all(np.isfinite(X)) and status == 0 and fail == -1
    prod_ms64 = f.products_ms()
    tf = C.c_double()
    A.capi.check(A.capi.lib().aslam_probe_dmma_tflops(dev.index, C.byref(tf)))
    step_ms = float(np.mean(ms))
    N, NS = n, 2 * n + 5
    # TF32 variant of the three sigma-point products (tcgen05 kind::tf32, accumulators in TMEM), separately toleranced:
    # same state and message, compared with the FP64 path of this run
    t32 = C.c_double()
    A.capi.check(A.capi.lib().aslam_probe_tf32_tflops(dev.index, C.byref(t32)))
    prod_flop = 2.0 * N * N * NS * 2  # P and M lower halves + Tc full = 2 full products
    tf32_rows = {}
    for mode, name in ((1, "tf32"), (3, "tf32x3")):
        f.set_precision(mode)
        tms, pms = [], []
        for it in range(3 + min(args.steps, 10)):
            f.load_state(X0, P0)
            f.step(0.01, -0.01, 0.002, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
            if it >= 3:
                tms.append(f.last_kernel_ms())
                pms.append(f.products_ms())
        Xt, Pt = f.get_state()
        st_t, fail_t = f.status()
        assert st_t == 0 and fail_t == -1
        issued = prod_flop * (1 if mode == 1 else 3)
        tf32_rows[name] = {
            "ms_per_step": float(np.mean(tms)), "products_ms": float(np.mean(pms)), "products_ms_fp64": prod_ms64,
            "relerr_X_vs_fp64": float(np.max(np.abs(Xt - X)) / np.max(np.abs(X))),
            "relerr_P_vs_fp64": float(np.max(np.abs(Pt - P)) / np.max(np.abs(P))),
            "roofline": {"bound": "tensor", "kernel": "tf32_gemm_nt_kernel<%d> (TMA + tcgen05.mma kind::tf32, TMEM)" % mode,
                         "unit": "TFLOP/s", "achieved": issued / (float(np.mean(pms)) * 1e-3) / 1e12, "peak": t32.value,
                         "frac": issued / (float(np.mean(pms)) * 1e-3) / 1e12 / t32.value if t32.value else None,
                         "note": "issued TF32 flop of the three products (x3 for the hi/lo split) over their "
                                 "CUDA-event time, FP64->hi/lo split kernels included",
                         "peak_source": "aslam_probe_tf32_tflops (kind::tf32 M128 N256 K8 from shared memory, this run)"}}
    f.close()
    executed = ((N ** 3) / 3 * 2 + 2 * N * N * NS + N * N * NS * 2 + N ** 3 * 2)  # 2 chol, P+M lower, Tc full, trsm+syrk
    reference_form = (N + 2) ** 3 / 3 + 3 * 2 * N * N * NS + 2 * N ** 3 * 4       # ~176 GF
    return {"workload": "cfg5: large-map UKF n=2051 (1024 landmarks), 4107 sigma points, FP64",
            "steps_per_s": 1e3 / step_ms, "ms_per_step": step_ms, "launches_per_step": launches,
            "roofline": {"bound": "tensor", "kernel": "dgemm_nt_kernel (TMA + DMMA.8x8x4) chain", "unit": "TFLOP/s",
                         "achieved": executed / (step_ms * 1e-3) / 1e12, "peak": tf.value,
                         "frac": executed / (step_ms * 1e-3) / 1e12 / tf.value if tf.value else None,
                         "executed_flop_per_step": executed, "reference_form_flop_per_step": reference_form,
                         "peak_source": "aslam_probe_dmma_tflops (DMMA chain, this run)"},
            "products_ms": prod_ms64, "tf32_variant": tf32_rows}


_REAL_STDOUT = None


def _claim_stdout():
    """Route fd 1 to stderr for the whole run and keep the real stdout for the ONE JSON line: libraries
    (NCCL prints its version banner to stdout on some boxes) must not add lines to it."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def _emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def laser_leg(args, torch, dev):
    """the laser front end. 65,536 synthetic LDS scans of the cfg3 map (one per filter of the batch);
    device-resident scans/s (CUDA events on the handle's stream), the same through host buffers (copy in, lists out),
    and the oracle on the host cores over a bounded sample. Algorithmic bytes per scan: 360 ranges in (1,440 B) +
    2 x 32 floats + 1 count out (260 B)."""
    import awesomeslam_b200 as A
    from oracle import oracle as O
    B = 65536
    lmx, lmy = A.map_cfg3()
    base = A.laser_scans(lmx, lmy, A.laser_poses(4096, seed=77, **A.replay.LASER_POSES_CFG3), seed=78)
    sc = np.tile(base, (B // base.shape[0], 1))
    lm = A.LandMarks(max_scans=B, device=dev.index)
    stream = torch.cuda.ExternalStream(lm.stream, device=dev)
    d_sc = torch.from_numpy(sc).to(dev)
    d_r = torch.empty((B, 32), dtype=torch.float32, device=dev)
    d_b = torch.empty((B, 32), dtype=torch.float32, device=dev)
    d_c = torch.empty((B,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)
    reps = 5
    with torch.cuda.stream(stream):
        for _ in range(2):
            lm.callback_device(d_sc.data_ptr(), B, d_r.data_ptr(), d_b.data_ptr(), d_c.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            lm.callback_device(d_sc.data_ptr(), B, d_r.data_ptr(), d_b.data_ptr(), d_c.data_ptr())
        e1.record()
        lm.sync()
        ms = e0.elapsed_time(e1) / reps
    # end to end through the host entry point, pinned host buffers: scans in, lists out, synchronous call
    h_sc = torch.from_numpy(sc).pin_memory()
    h_r = torch.zeros((B, 32), dtype=torch.float32).pin_memory()
    h_b = torch.zeros((B, 32), dtype=torch.float32).pin_memory()
    h_c = torch.zeros((B,), dtype=torch.int32).pin_memory()
    lm.callback_into(h_sc.numpy(), h_r.numpy(), h_b.numpy(), h_c.numpy())  # warm-up
    t0 = time.perf_counter()
    for _ in range(3):
        lm.callback_into(h_sc.numpy(), h_r.numpy(), h_b.numpy(), h_c.numpy())
    host_s = (time.perf_counter() - t0) / 3
    rg, bg, cnt = h_r.numpy(), h_b.numpy(), h_c.numpy()
    sample = sc[:4096]
    from oracle import ref as R
    if R.available("laser"):  # the reference's own sensor_landmark.cpp, compiled unmodified (oracle/_ref)
        org, obg, ocnt, sec = R.laser_extract(sample, nthreads=host_threads())
        lkind, lcores = "reference", host_threads()
    else:
        org, obg, ocnt, sec = O.laser_extract(sample)
        lkind, lcores = "port", O.num_threads()
    assert np.array_equal(cnt[:4096], ocnt) and np.array_equal(d_c.cpu().numpy(), cnt) and int(cnt.min()) >= 3
    same = all(np.array_equal(rg[i, :ocnt[i]], org[i, :ocnt[i]]) for i in range(0, 4096, 64))
    lm.close()
    return {"workload": "laser front end: 65,536 LDS scans (360 rays) of the cfg3 map -> (range, bearing) lists",
            "scans_per_s": B / (ms * 1e-3), "ms_per_batch": ms, "landmarks_per_scan": float(cnt.mean()),
            "hbm_gbs": B * (1440 + 260) / (ms * 1e-3) / 1e9,
            "e2e_scans_per_s": B / host_s, "e2e_api": "aslam_laser_extract (pinned host buffers, chunked: copy-in of chunk k+1 under the kernels of chunk k)",
            "e2e_h2d_bytes": B * 1440, "e2e_d2h_bytes": B * 260,
            "cpu_baseline": {"value": 4096 / sec, "unit": "scans/s", "cores": lcores, "kind": lkind,
                             "sample": "4,096 of the 65,536 scans, OpenMP over scans"},
            "matches_oracle_on_sample": bool(same)}


def main():
    args = parse()
    _claim_stdout()
    if args.impl == "reference":
        reference_arm(args)
        return
    import torch
    import torch.distributed as dist
    import awesomeslam_b200 as A
    A.capi.lib()  # fails loudly when the CUDA library is missing: there is no fallback path
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU path)")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    hbm_peak, peak_src = measured_peaks()

    res = batched_leg(args, torch, dist, rank, world, dev)
    B, K, W = args.batch, args.steps, args.warmup
    updates = B * world * FUSE * K
    value = updates / (res["resident"]["ms"] * 1e-3)
    e2e = updates / (res["e2e"]["ms"] * 1e-3)

    fp64 = None
    if rank == 0:
        import ctypes as C
        tf = C.c_double()
        A.capi.check(A.capi.lib().aslam_probe_fp64_tflops(local, C.byref(tf)))
        fp64 = tf.value
    large = None
    if not args.no_large:
        large = large_leg(args, torch, dist, rank, world, dev, hbm_peak, peak_src)
    # everything that involves more than one rank is done: the other ranks leave now instead of spinning in a barrier
    # while rank 0 runs its single-GPU legs and the CPU baselines
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    ukf_large = None
    if not args.no_large:
        ukf_large = ukf_large_leg(args, torch, dev)
        if not args.no_cpu:
            ukf_large["cpu_baseline"] = ukf_large_cpu_leg()
            if large is not None and world == 1:
                large["cpu_baseline"] = large_cpu_leg()
    laser = None
    ukf_batched = ukf_batched35 = None
    if not args.no_large:
        laser = laser_leg(args, torch, dev)
        ukf_batched = ukf_batched_leg(args, torch, dev)
        ukf_batched["max_state_dim_20"] = ukf_batched_leg(args, torch, dev, shape="cfg2_cap20")
        ukf_batched35 = ukf_batched_leg(args, torch, dev, B=65536, shape="cfg3")

    per_gpu_ups = value / world
    line = {
        "metric": "ekf_slam_filter_updates_per_s", "value": value, "unit": "updates/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": res["resident"]["ms"] / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "cfg3: 65,536 EKF-SLAM filters x 16 landmarks per GPU (n=35), Monte Carlo seeds",
                   "filters_per_gpu": B, "messages_per_step": FUSE, "updates_per_step": B * world * FUSE,
                   "l2": "inputs larger than L2 (covariances %.0f MB + messages per GPU)" % (B * 36 * 36 * 8 / 1e6),
                   "parallelism": "batch-sharded, no collective" if world > 1 else "1 GPU"},
        "e2e": {"value": e2e, "unit": "updates/s", "h2d_bytes_per_step": res["h2d_bytes_per_step"],
                "d2h_bytes_per_step": res["d2h_bytes_per_step"], "ms_per_step": res["e2e"]["ms"] / K,
                "api": "aslam_filter_run_host (pinned host messages, copy overlapped with the previous kernel) + "
                       "aslam_filter_read_state_async/read_wait every step"},
        "gpu_launches": res["resident"]["launches"],
        "clocks": res["resident"]["clocks"],
        "parity": res.get("parity"),
        "roofline_batched": {
            "kernel": "ekf_batched_kernel<36,2>", "bound": "fp64",
            "achieved": per_gpu_ups * FLOP_PER_UPDATE / 1e12, "peak": fp64, "unit": "TFLOP/s",
            "frac": (per_gpu_ups * FLOP_PER_UPDATE / 1e12 / fp64) if fp64 else None,
            "flop_convention": "6 n^3 = 257,250 per update (SURVEY 8d); the kernel executes ~n^3 fma",
            "achieved_executed": per_gpu_ups * 2.0 * 35 ** 3 / 1e12,
            "frac_executed": (per_gpu_ups * 2.0 * 35 ** 3 / 1e12 / fp64) if fp64 else None,
            "pipe_fp64_active_pct_ncu": _ncu_batched_pipe(),
            "hbm_frac_if_streamed": per_gpu_ups * BYTES_PER_UPDATE / 1e9 / hbm_peak,
            "peak_source": "aslam_probe_fp64_tflops (DFMA chain, this run)"},
    }
    if large:
        line["roofline"] = large["roofline"]
        line["large_map"] = {k: v for k, v in large.items() if k != "roofline"}
    else:
        line["roofline"] = None
    if ukf_large:
        line["large_map_ukf"] = ukf_large
    if laser is not None:
        line["laser_front"] = laser
    if ukf_batched is not None:
        line["batched_ukf"] = ukf_batched
    if ukf_batched35 is not None:
        line["batched_ukf_cfg3_shape"] = ukf_batched35
    if not args.no_cpu:
        cpu_filters, cpu_msgs = (1024 if host_threads() < 16 else 4096), 240
        line["cpu_baseline"] = cpu_leg(cpu_filters, cpu_msgs)
        line["speedup_e2e_vs_cpu"] = e2e / line["cpu_baseline"]["value"]
    _emit(line)


if __name__ == "__main__":
    main()
