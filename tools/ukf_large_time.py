"""cfg5 step time (CUDA events inside the library): python tools/ukf_large_time.py [side] [reps]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A
side = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
tx, ty, X0, G, d = A.large_map_state(side, pitch=100.0 / side)
P0 = A.dense_P(G, d)
rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
f = A.LargeFilter("ukf", X0.size)
ms = []
for it in range(3 + reps):
    f.load_state(X0, P0)
    f.step(0.01, -0.01, 0.002, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
    if it >= 3:
        ms.append(f.last_kernel_ms())
st = f.status()
print("n=%d: %.3f ms per step (min %.3f), products %.3f ms, status %s" % (X0.size, np.mean(ms), np.min(ms), f.products_ms(), st))
f.close()
