"""Dump the state of a small batched-UKF replay (debug aid: the result must not depend on the CTA width).
usage: python tools/ukf_dump.py out.npz [B] [T]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A  # noqa: E402

B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
T = int(sys.argv[3]) if len(sys.argv) > 3 else 80
lmx, lmy = A.map_cfg1()
rp = A.Replay(lmx, lmy, B, T, seed0=0x0CF20000)
f = A.BatchedFilter("ukf", B, 0, max_state_dim=30)
snaps = {}
for t0 in range(0, T, 8):
    sl = slice(t0, t0 + 8)
    f.run_host(8, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl], rp.range[sl], rp.bearing[sl], len(lmx))
    st = f.get_state()
    snaps["X%d" % (t0 + 8)] = st["X"]
    snaps["S%d" % (t0 + 8)] = st["status"]
np.savez(sys.argv[1], **snaps)
print("dumped", sys.argv[1])
