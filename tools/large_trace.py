"""Timeline of the large-map EKF step chain (front / solve / streaming update) from %globaltimer stamps.
   ASLAM_LARGE_TRACE=1 python tools/large_trace.py [L_side] [m] [rows: begin end]"""
import os, sys
import numpy as np
os.environ["ASLAM_LARGE_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A
from awesomeslam_b200 import capi

L_side = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m = int(sys.argv[2]) if len(sys.argv) > 2 else 8
tx, ty, X0, G, d = A.large_map_state(L_side)
P0 = A.dense_P(G, d)
n = X0.size
f = A.LargeFilter("ekf", n)
f.load_state_colmajor(X0, P0)
rng, brg, idx = A.large_observations(tx, ty, (0, 0, 0), m=m)
steps = 40
for _ in range(steps):
    f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
f.sync()
tr = np.zeros((64, 16), np.uint64)
capi.check(f._lib.aslam_large_trace(f._h, capi.ptr(tr)))
tr = tr.astype(np.int64)
rows = []
for k in range(steps - 12, steps):
    a, b = tr[k % 64], tr[(k + 1) % 64]
    rows.append(dict(front=(a[1] - a[0]) / 1e3, solve=(a[3] - a[2]) / 1e3, front_to_solve=(a[2] - a[1]) / 1e3,
                     upd_in_after_solve_in=(a[6] - a[2]) / 1e3, solve_out_to_upd_go=(a[4] - a[3]) / 1e3,
                     update=(a[5] - a[4]) / 1e3, upd_out_to_next_solve_in=(b[2] - a[5]) / 1e3,
                     next_front_start_after_solve_out=(b[0] - a[3]) / 1e3, period=(b[2] - a[2]) / 1e3,
                     s_gathered=(a[8] - a[2]) / 1e3, s_psub=(a[9] - a[8]) / 1e3, s_built=(a[10] - a[9]) / 1e3,
                     s_factor=(a[11] - a[10]) / 1e3, s_kx=(a[12] - a[11]) / 1e3, s_cta0_to_last_out=(a[3] - a[12]) / 1e3))
keys = list(rows[0].keys())
print("us per step (median of %d steps):" % len(rows))
for k in keys:
    print("  %-34s %8.1f" % (k, float(np.median([r[k] for r in rows]))))
