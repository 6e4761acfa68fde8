"""Run-to-run reproducibility of the multi-stream / shared-memory paths (a cheap detector of ordering bugs):
batched EKF in two launch splits at a large batch, large-map UKF step (side streams: Cholesky lookahead, ragged-edge
GEMM rows) three times from the same state, large-map EKF steps on two handles."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A  # noqa: E402

ok = True
# ---- batched EKF
lmx, lmy = A.map_cfg3()
B, T = 16384, 48
rp = A.Replay(lmx, lmy, B, T, seed0=A.replay.SEED_CFG3)
outs = []
for split in (48, 8):
    f = A.BatchedFilter("ekf", B, 0, max_state_dim=36)
    for t0 in range(0, T, split):
        sl = slice(t0, t0 + split)
        f.run_host(split, rp.x[sl], rp.y[sl], rp.yaw[sl], rp.vx[sl], rp.az[sl], rp.dt[sl], rp.range[sl], rp.bearing[sl], 16)
    st = f.get_state()
    outs.append(st)
    f.close()
same = all(np.array_equal(outs[0][k], outs[1][k]) for k in ("X", "P", "N", "status"))
print("batched EKF %d filters x %d messages, 1 launch vs 6 launches: %s" % (B, T, "identical" if same else "DIFFERENT"))
ok &= same
# ---- large UKF
side = 32
tx, ty, X0, G, d = A.large_map_state(side, pitch=100.0 / side)
P0 = A.dense_P(G, d)
rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
f = A.LargeFilter("ukf", X0.size)
res = []
for it in range(3):
    f.load_state(X0, P0)
    f.step(0.01, -0.01, 0.002, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    res.append(f.get_state())
f.close()
same = all(np.array_equal(res[0][0], r[0]) and np.array_equal(res[0][1], r[1]) for r in res[1:])
print("large UKF n=%d, the same step three times: %s" % (X0.size, "identical" if same else "DIFFERENT"))
ok &= same
# ---- large EKF
tx, ty, X0, G, d = A.large_map_state(64)
P0 = A.dense_P(G, d)
rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
res = []
for it in range(2):
    f = A.LargeFilter("ekf", X0.size)
    f.load_state_colmajor(X0, P0)
    for k in range(6):
        f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    res.append(f.get_state())
    f.close()
same = np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])
print("large EKF n=%d, six steps on two handles: %s" % (X0.size, "identical" if same else "DIFFERENT"))
ok &= same
sys.exit(0 if ok else 1)
