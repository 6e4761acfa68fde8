"""Timeline of the row-sharded large-map EKF step with the fused exchange on G GPUs of one box (handles of ONE process,
peer access): %globaltimer stamps of every shard.   python tools/large_trace_peer.py [G] [L_side] [m]"""
import os, sys, time
import numpy as np
TRACE = os.environ.get("NOTRACE") is None
if TRACE:
    os.environ["ASLAM_LARGE_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A
from awesomeslam_b200 import capi

Gn = int(sys.argv[1]) if len(sys.argv) > 1 else 2
L_side = int(sys.argv[2]) if len(sys.argv) > 2 else 64
m = int(sys.argv[3]) if len(sys.argv) > 3 else 8
tx, ty, X0, G, d = A.large_map_state(L_side)
P0 = A.dense_P(G, d)
n = X0.size
edges = [0] + [max(3, (n * r) // Gn) for r in range(1, Gn)] + [n]
shards = [A.LargeFilter("ekf", n, r, row_begin=edges[r], row_end=edges[r + 1]) for r in range(Gn)]
for f in shards:
    f.load_state_colmajor(X0, P0)
blocks = [f.peer_export()[1] for f in shards]
for r, f in enumerate(shards):
    f.peer_attach(Gn, r, blocks=blocks)
rng, brg, idx = A.large_observations(tx, ty, (0, 0, 0), m=m)
steps = 40
import torch
def run(nsteps):
    for _ in range(nsteps):
        for f in shards:
            f.step_fused(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
run(10)
for f in shards:
    f.sync()
evs = []
for r, f in enumerate(shards):
    st = torch.cuda.ExternalStream(f.stream, device=torch.device("cuda", r))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    evs.append((st, e0, e1))
    with torch.cuda.device(r):
        e0.record(st)
t0 = time.perf_counter()
run(steps)
for r, f in enumerate(shards):
    with torch.cuda.device(r):
        evs[r][2].record(evs[r][0])
for f in shards:
    f.sync()
wall = (time.perf_counter() - t0) / steps * 1e3
dev_ms = [e0.elapsed_time(e1) / steps for (_, e0, e1) in evs]
print("G=%d n=%d: %s ms per step by CUDA events per shard; %.4f ms host wall clock (%d steps, one host thread enqueues all shards)"
      % (Gn, n, ["%.4f" % v for v in dev_ms], wall, steps))
if not TRACE:
    for f in shards:
        assert f.status()[0] == 0
        f.close()
    sys.exit(0)
for r, f in enumerate(shards):
    assert f.status()[0] == 0
    tr = np.zeros((64, 16), np.uint64)
    capi.check(f._lib.aslam_large_trace(f._h, capi.ptr(tr)))
    tr = tr.astype(np.int64)
    rows = []
    for k in range(steps - 12, steps - 1):
        a, b = tr[k % 64], tr[(k + 1) % 64]
        rows.append(dict(front=(a[1] - a[0]) / 1e3, front_to_solve_in=(a[2] - a[1]) / 1e3, solve=(a[3] - a[2]) / 1e3,
                         s_gathered=(a[8] - a[2]) / 1e3, s_psub=(a[9] - a[8]) / 1e3, s_built=(a[10] - a[9]) / 1e3,
                         s_factor=(a[13] - a[10]) / 1e3, s_sync=(a[11] - a[13]) / 1e3, s_kx=(a[12] - a[11]) / 1e3,
                         solve_out_to_upd_go=(a[4] - a[3]) / 1e3, update=(a[5] - a[4]) / 1e3,
                         upd_out_to_next_solve_in=(b[2] - a[5]) / 1e3, period=(b[2] - a[2]) / 1e3))
    print("shard %d rows [%d, %d): us (median of %d steps)" % (r, edges[r], edges[r + 1], len(rows)))
    print("   " + "  ".join("%s %.1f" % (k, float(np.median([q[k] for q in rows]))) for k in rows[0]))
for f in shards:
    f.close()
