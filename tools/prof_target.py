"""Short single-GPU programs for ncu captures (B200_PROFILING.md): a handful of launches of one kernel.

  python tools/prof_target.py batched [B] [msgs]   cfg3 batched EKF kernel, B filters at N = 35
  python tools/prof_target.py ukf [B] [msgs]       batched UKF kernel, cfg3 shape (n = 35); ukf19: the cfg2 shape (n = 19)
  python tools/prof_target.py large [side] [m]     cfg4 large-map EKF step (side^2 landmarks)
  python tools/prof_target.py ukf_large [side] [p] cfg5 large-map UKF step (p = 0 FP64, 1 TF32, 3 3xTF32 products)
  python tools/prof_target.py laser [B]            laser front end, B scans of the cfg3 map
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A  # noqa: E402


def batched(kind, B, msgs, small=False):
    lmx, lmy = A.map_cfg1() if small else A.map_cfg3()
    L = len(lmx)
    warm = 24
    rp = A.Replay(lmx, lmy, B, warm + 4 * msgs, seed0=A.replay.SEED_CFG3)
    f = A.BatchedFilter(kind, B, 0, max_state_dim=30 if small else 36)
    a = [rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing]
    f.run_host(warm, *[v[:warm] for v in a], L)
    f.sync()
    for k in range(4):
        s = slice(warm + k * msgs, warm + (k + 1) * msgs)
        f.run_host(msgs, *[v[s] for v in a], L)
    f.sync()
    st = f.get_state(want_P=False, want_assoc=False)
    print(kind, "N", int(st["N"].min()), int(st["N"].max()), "status", int(st["status"].max()))


def large(side, m):
    tx, ty, X0, G, d = A.large_map_state(side)
    P0 = A.dense_P(G, d)
    f = A.LargeFilter("ekf", X0.size)
    f.load_state_colmajor(X0, P0)
    rng, brg, idx = A.large_observations(tx, ty, (0, 0, 0), m=m)
    for _ in range(6):
        f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    f.sync()
    print("large n", X0.size)


def ukf_large(side, precision=0):
    tx, ty, X0, G, d = A.large_map_state(side, pitch=100.0 / side)
    P0 = A.dense_P(G, d)
    f = A.LargeFilter("ukf", X0.size)
    if precision:
        f.set_precision(precision)
    f.load_state_colmajor(X0, P0)
    rng, brg, idx = A.large_observations(tx, ty, (0, 0, 0), m=8)
    for _ in range(3):
        f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    f.sync()
    print("ukf_large n", X0.size)


def laser(B):
    lmx, lmy = A.map_cfg3()
    base = A.laser_scans(lmx, lmy, A.laser_poses(4096, seed=77, **A.replay.LASER_POSES_CFG3), seed=78)
    sc = np.tile(base, (max(B // 4096, 1), 1))
    import torch
    n = sc.shape[0]
    lm = A.LandMarks(max_scans=n)
    dev = torch.device("cuda", 0)
    d_sc = torch.from_numpy(sc).to(dev)
    d_r = torch.empty((n, 32), dtype=torch.float32, device=dev)
    d_b = torch.empty((n, 32), dtype=torch.float32, device=dev)
    d_c = torch.empty((n,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)
    for _ in range(3):  # the device-resident path: one classification + one fit launch for the whole batch
        lm.callback_device(d_sc.data_ptr(), n, d_r.data_ptr(), d_b.data_ptr(), d_c.data_ptr())
    lm.sync()
    print("laser scans", n, "landmarks/scan", float(d_c.float().mean().item()))


if __name__ == "__main__":
    mode = sys.argv[1]
    arg = [int(v) for v in sys.argv[2:]]
    if mode == "laser":
        laser(arg[0] if arg else 65536)
    elif mode in ("batched", "ukf", "ukf19"):
        batched("ekf" if mode == "batched" else "ukf", arg[0] if arg else 65536, arg[1] if len(arg) > 1 else 8,
                small=mode == "ukf19")
    elif mode == "large":
        large(arg[0] if arg else 64, arg[1] if len(arg) > 1 else 8)
    elif mode == "ukf_large":
        ukf_large(arg[0] if arg else 32, arg[1] if len(arg) > 1 else 0)
