"""TEMP: time the cfg3 batched EKF kernel with parts switched off (ASLAM_K1_DBG bits) after a normal warm-up."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import awesomeslam_b200 as A
B, FUSE, WARM, K = 65536, 8, 24, 6
dev = torch.device("cuda", 0)
lmx, lmy = A.map_cfg3()
rp = A.Replay(lmx, lmy, B, WARM + (K + 2) * FUSE, seed0=A.replay.SEED_CFG3)
dev_t = [torch.from_numpy(a).to(dev) for a in [rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing]]
def run(f, msg0, n):
    p = [t.data_ptr() + msg0 * t[0].numel() * t.element_size() for t in dev_t]
    f.run_device(n, *p, 16)
for dbg in [int(a) for a in sys.argv[1:]] or [0]:
    os.environ["ASLAM_K1_DBG"] = "0"
    f = A.BatchedFilter("ekf", B, 0, max_state_dim=36)
    stream = torch.cuda.ExternalStream(f.stream, device=dev)
    run(f, 0, WARM); f.sync()
    os.environ["ASLAM_K1_DBG"] = str(dbg)
    run(f, WARM, FUSE); run(f, WARM + FUSE, FUSE)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record()
        for k in range(K):
            run(f, WARM + (2 + k) * FUSE - 2 * FUSE + 2 * FUSE, FUSE)
        e1.record()
    f.sync()
    ms = e0.elapsed_time(e1) / K
    print("dbg %2d: %.3f ms per launch of %d messages -> %.1f M updates/s" % (dbg, ms, FUSE, B * FUSE / ms / 1e3))
    f.close()
