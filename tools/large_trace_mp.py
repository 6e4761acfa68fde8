"""Timeline of the row-sharded large-map EKF step with the fused exchange, ONE PROCESS PER GPU (CUDA IPC mappings, gloo
for the handle exchange and the barriers: the set-up bench.py times under torchrun, without NCCL).
   python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29520 tools/large_trace_mp.py [steps]"""
import os, sys, time
import numpy as np
if os.environ.get("NOTRACE") is None:
    os.environ["ASLAM_LARGE_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import awesomeslam_b200 as A
from awesomeslam_b200 import capi

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
if os.environ.get("BACKEND") == "nccl":
    dist.init_process_group("nccl", device_id=dev)
else:
    dist.init_process_group("gloo")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
tx, ty, X0, G, d = A.large_map_state(int(os.environ.get('SIDE', '64')))
P0 = A.dense_P(G, d)
n = X0.size
rb = 0 if rank == 0 else max(3, (n * rank) // world)
re = n if rank == world - 1 else max(3, (n * (rank + 1)) // world)
f = A.LargeFilter("ekf", n, dev.index, row_begin=rb, row_end=re)
f.load_state_colmajor(X0, P0)
hd, _ = f.peer_export()
handles = [None] * world
dist.all_gather_object(handles, hd)
f.peer_attach(world, rank, ipc_handles=handles)
dist.barrier()
rng, brg, idx = A.large_observations(tx, ty, (0, 0, 0), m=8)
def run(k):
    for _ in range(k):
        f.step_fused(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg, want_assoc=False)
if os.environ.get("PRE_NCCL_FORM"):  # what bench.py does first on the same handle: steps of the NCCL all-reduce form
    class _DevArray:
        def __init__(self, ptr, count):
            self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 2}
    stream0 = torch.cuda.ExternalStream(f.stream, device=dev)
    with torch.cuda.stream(stream0):
        exch = None
        for _ in range(20):
            f.step_phase1(0.0, 0.0, 0.0, 0.05, 0.1, 1.0, rng, brg)
            if exch is None:
                p_, c_ = f.exchange_buffer()
                exch = torch.as_tensor(_DevArray(p_, c_), device=dev)
            dist.all_reduce(exch)
            f.step_phase2(want_assoc=False)
        f.sync()
run(10)
f.sync()
dist.barrier()
st = torch.cuda.ExternalStream(f.stream, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
run(steps)
e1.record(st)
f.sync()
ms = e0.elapsed_time(e1) / steps
assert f.status()[0] == 0
out = "rank %d rows [%d, %d): %.4f ms per step" % (rank, rb, re, ms)
if os.environ.get("NOTRACE") is None:
    tr = np.zeros((64, 16), np.uint64)
    capi.check(f._lib.aslam_large_trace(f._h, capi.ptr(tr)))
    tr = tr.astype(np.int64)
    rows = []
    for k in range(10 + steps - 12, 10 + steps - 1):
        a, b = tr[k % 64], tr[(k + 1) % 64]
        rows.append(dict(front_to_solve_in=(a[2] - a[1]) / 1e3, solve=(a[3] - a[2]) / 1e3, s_gathered=(a[8] - a[2]) / 1e3,
                         s_psub=(a[9] - a[8]) / 1e3, s_built=(a[10] - a[9]) / 1e3, s_factor=(a[13] - a[10]) / 1e3,
                         s_kx=(a[12] - a[11]) / 1e3, solve_out_to_upd_go=(a[4] - a[3]) / 1e3, update=(a[5] - a[4]) / 1e3,
                         upd_out_to_next_solve_in=(b[2] - a[5]) / 1e3, next_front_end_after_solve_out=(b[1] - a[3]) / 1e3,
                         period=(b[2] - a[2]) / 1e3))
    out += "\n   " + "  ".join("%s %.1f" % (k, float(np.median([q[k] for q in rows]))) for k in rows[0])
for r in range(world):
    if r == rank:
        print(out, flush=True)
    dist.barrier()
f.close()
dist.destroy_process_group()
