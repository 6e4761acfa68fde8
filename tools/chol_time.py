"""Time the blocked Cholesky pieces on one GPU: python tools/chol_time.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A

lib = A.capi.lib()
st = torch.cuda.current_stream().cuda_stream
for n in (64, 128, 512, 2053):
    ld = ((n + 15) // 16) * 16
    g = torch.randn(n, n, dtype=torch.float64, device="cuda")
    M = g @ g.T + n * torch.eye(n, dtype=torch.float64, device="cuda")
    buf = torch.zeros(n, ld, dtype=torch.float64, device="cuda")
    inv = torch.zeros(((n + 63) // 64) * 64 * 64, dtype=torch.float64, device="cuda")
    fail = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    reps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for r in range(reps + 3):
        buf[:, :n] = M
        torch.cuda.synchronize()
        e0.record()
        A.capi.check(lib.aslam_dense_cholesky(st, buf.data_ptr(), ld, n, inv.data_ptr(), fail.data_ptr()))
        e1.record()
        torch.cuda.synchronize()
        if r >= 3:
            tot += e0.elapsed_time(e1)
    print("cholesky n=%d: %.1f us  (%d panels -> %.1f us/panel step)" % (n, 1e3 * tot / reps, (n + 63) // 64, 1e3 * tot / reps / ((n + 63) // 64)))
