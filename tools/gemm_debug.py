"""Debug helper for the TMA + DMMA GEMM: which K slabs / tiles are wrong."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import awesomeslam_b200 as A  # noqa: E402

lib = A.capi.lib()


def gemm(a, b, c, alpha=1.0, beta=0.0, lower=0):
    M, K = a.shape
    N = b.shape[0]
    A.capi.check(lib.aslam_dense_gemm_nt(None, M, N, K, a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0),
                                         c.data_ptr(), c.stride(0), alpha, beta, lower))
    torch.cuda.synchronize()


def padded(rows, cols, fill=None):
    ld = (cols + 15) // 16 * 16
    buf = torch.zeros((rows, ld), dtype=torch.float64, device="cuda")
    v = buf[:, :cols]
    if fill is None:
        v.copy_(torch.randn(rows, cols, dtype=torch.float64, device="cuda"))
    else:
        v.fill_(fill)
    return v


for (M, N, K) in [(128, 64, 64), (128, 64, 80), (128, 64, 128), (128, 64, 1030), (515, 259, 1030), (256, 256, 4107)]:
    a, b, c = padded(M, K), padded(N, K), padded(M, N, 0.0)
    gemm(a, b, c)
    ref = a @ b.T
    err = (c - ref).abs()
    print("M %d N %d K %d  max err %.3e  wrong entries %d / %d" % (M, N, K, err.max().item(), int((err > 1e-9).sum()), M * N))
    if err.max().item() > 1e-9:
        KT = (K + 15) // 16
        bad = []
        for s in range(KT):
            a2 = padded(M, K, 0.0)
            a2[:, s * 16:(s + 1) * 16] = a[:, s * 16:(s + 1) * 16]
            c2 = padded(M, N, 0.0)
            gemm(a2, b, c2)
            e2 = (c2 - a2 @ b.T).abs().max().item()
            if e2 > 1e-9:
                bad.append((s, "%.2e" % e2))
        print("   bad slabs:", bad[:40])
        rows = (err > 1e-9).any(dim=1).nonzero().flatten().tolist()
        cols = (err > 1e-9).any(dim=0).nonzero().flatten().tolist()
        print("   bad rows:", rows[:20], "... count", len(rows), " bad cols:", cols[:20], "... count", len(cols))
        # repeatability
        c3 = padded(M, N, 0.0)
        gemm(a, b, c3)
        print("   repeat differs from first run:", (c3 - c).abs().max().item())
