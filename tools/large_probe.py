"""Times the large-map EKF step (cfg4) on one GPU: whole step and the streaming update kernel."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A

L_side = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m = int(sys.argv[2]) if len(sys.argv) > 2 else 8
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 30
tx, ty, X0, G, d = A.large_map_state(L_side)
P0 = A.dense_P(G, d)
n = X0.size
f = A.LargeFilter("ekf", n)
f.load_state_colmajor(X0, P0)
rng, brg, idx = A.large_observations(tx, ty, (0, 0, 0), m=m)
for _ in range(5):
    f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
f.sync()
t0 = time.perf_counter()
for _ in range(steps):
    f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
f.sync()
dt = (time.perf_counter() - t0) / steps
ks = []
f.set_timing(True)
for _ in range(steps):
    f.step(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    ks.append(f.last_kernel_ms())
bytes_ = 2 * 8 * n * n
k = float(np.mean(ks))
print("n=%d q=%d step %.3f ms (%.0f steps/s)  update kernel %.3f ms (min %.3f)  %.0f GB/s algorithmic" %
      (n, 3 + 2 * m, dt * 1e3, 1 / dt, k, min(ks), bytes_ / (k * 1e-3) / 1e9))
