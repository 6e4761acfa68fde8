"""world_size-2 tests of the multi-GPU host logic on CPU (torch.distributed, gloo backend, 127.0.0.1).

(a) Batched filters shard by global filter id with no collective: two ranks, each replaying its own shard
    (generated from first_filter = rank * per_gpu), reproduce the unsharded batch bit for bit.
(b) The large-map EKF shards P by row blocks with ONE exchange per step: every rank fills the [2q][n] exchange
    buffer for the rows it owns (awesomeslam_b200.sharding.exchange_contribution pins the layout the CUDA gather
    kernel writes), the buffers are summed (gloo here, NCCL on the GPUs), and each rank finishes its own rows
    locally. The assembled result must equal the oracle's unsharded step (golden vector large_ekf_l6.npz).
The arithmetic here is numpy/oracle on purpose: these tests cover partitioning and the exchange contract, the
CUDA kernels are covered by the -m gpu tests (1 GPU, and 2 ranks when the box has them).
"""
import math
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(fn, world=2):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(fn, args=(world, port), nprocs=world, join=True)


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    return dist


# ------------------------------------------------------------------ (a) batch sharding
def _batch_worker(rank, world, port):
    dist = _init(rank, world, port)
    import torch
    import awesomeslam_b200 as A
    from awesomeslam_b200 import sharding
    from oracle import oracle as O
    per, T, L = 3, 30, 16
    lo, hi = sharding.filter_range(per, rank, world)
    lmx, lmy = A.map_cfg3()
    rp = A.Replay(lmx, lmy, per, T, seed0=A.replay.SEED_CFG3, first_filter=lo)
    cfg = O.default_config(max_state_dim=36)
    _, X, P, N = O.run_batch("ekf", cfg, per, T, L, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing, 1, 36)
    mine = torch.from_numpy(np.concatenate([X.ravel(), P.ravel(), N.astype(np.float64)]))
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)  # result gather only: the data path has no collective
    if rank == 0:
        g = A.Replay(lmx, lmy, per * world, T, seed0=A.replay.SEED_CFG3)
        _, Xg, Pg, Ng = O.run_batch("ekf", cfg, per * world, T, L, g.x, g.y, g.yaw, g.vx, g.az, g.dt, g.range,
                                    g.bearing, 1, 36)
        for r in range(world):
            a, b = sharding.filter_range(per, r, world)
            ref = np.concatenate([Xg[a:b].ravel(), Pg[a:b].ravel(), Ng[a:b].astype(np.float64)])
            assert np.array_equal(parts[r].numpy(), ref), "shard %d differs from the unsharded batch" % r
        assert np.all(Ng == 35)
    dist.barrier()
    dist.destroy_process_group()


def test_batch_shards_equal_unsharded_batch():
    _spawn(_batch_worker)


# ------------------------------------------------------------------ (b) row-block sharding
def _h_rows(X, lm_slots):
    """Rows of H in use (ekf.cpp:117-134 with the float hyp/dist casts): pose rows, then (range, bearing) of each
    landmark slot in ascending order. Returns rows[], hcols[], hvals[]."""
    f32 = np.float32
    rows, hcols, hvals = [0, 1, 2], [[0], [1], [2]], [[1.0], [1.0], [1.0]]
    for i in sorted(lm_slots):
        dx, dy = X[3 + i] - X[0], X[4 + i] - X[1]
        hyp = f32(dx * dx + dy * dy)
        dist = float(f32(np.sqrt(hyp)))
        hyp = float(hyp)
        a, c = -dx / dist, -dy / dist       # H(3+i,0), H(3+i,1)
        b, d = dy / hyp, -dx / hyp          # H(4+i,0), H(4+i,1)
        rows += [3 + i, 4 + i]
        hcols += [[0, 1, 3 + i, 4 + i], [0, 1, 2, 3 + i, 4 + i]]
        hvals += [[a, c, -a, -c], [b, d, -1.0, -b, -d]]
    return rows, hcols, hvals


def _rows_worker(rank, world, port):
    dist = _init(rank, world, port)
    import torch
    from awesomeslam_b200 import sharding
    from oracle import np_restatement as R
    g = np.load(os.path.join(GOLD, "large_ekf_l6.npz"))
    X0, P0 = g["X0"], g["P0"]
    n = X0.size
    ox, oy, yaw, vx, az, dt = [float(v) for v in g["odom"]]
    rb, re = sharding.row_block(n, rank, world)
    assert (rb == 0) == (rank == 0) and re - rb >= 3
    Pl = P0[rb:re].copy()  # this rank's rows (P0 is symmetric, so row/col-major does not matter for the input)
    # --- front (replicated on every rank: O(n) work on X only)
    f32 = np.float32
    dth, r = float(f32(az * float(f32(dt)))), float(f32(vx / az))
    z2 = float(f32(yaw))
    a00 = r * (-math.cos(z2) + math.cos(z2 + dth))
    a10 = r * (-math.sin(z2) + math.sin(z2 + dth))
    X = R.state_transition(n, X0, f32(vx), f32(az), f32(dt))
    X[2] = float(R.normalize_angle(X[2]))
    # --- predict P = A P A' + Q on the local rows (rows 0,1 live on rank 0; columns 0,1 are local everywhere)
    kq = float(R.EKF_KQ)
    if rank == 0:
        r0 = Pl[0].copy()
        Pl[0] = a00 * r0
        Pl[1] = a10 * r0 + Pl[1]
    c0 = Pl[:, 0].copy()
    Pl[:, 0] = c0 * a00
    Pl[:, 1] = c0 * a10 + Pl[:, 1]
    if rank == 0:
        Pl[0, 0] += kq
        Pl[1, 1] += kq
        Pl[2, 2] += kq
    # --- the one exchange
    slots = [int(s) for s in g["assoc"]]
    rows, hcols, hvals = _h_rows(X, slots)
    q = len(rows)
    E = sharding.exchange_contribution(Pl, rb, rows, hcols, hvals, n)
    owners = np.zeros((2 * q, n))
    owners[E != 0] = 1
    Et, Ot = torch.from_numpy(E), torch.from_numpy(owners)
    sharding.exchange_sum(Et)
    sharding.exchange_sum(Ot)
    assert Ot.max().item() <= 1.0, "an element of the exchange buffer has two owners"
    E = Et.numpy()
    # --- phase 2, local: S, K for the owned rows, X, P_local -= K_local (H_o P)
    HP = np.zeros((q, n))
    for k in range(q):
        kk = {c: i for i, c in enumerate(rows)}
        for c, v in zip(hcols[k], hvals[k]):
            HP[k] += v * E[kk[c]]
    S = np.zeros((q, q))
    for b in range(q):
        for c, v in zip(hcols[b], hvals[b]):
            S[:, b] += HP[:, c] * v
    rdiag = np.array([float(R.EKF_KR)] * 3 + [float(R.UKF_KR)] * (q - 3))
    S += np.diag(rdiag)
    PHt = E[q:].T                      # [n][q], complete after the sum
    K = PHt @ np.linalg.inv(S)
    # innovation
    Zfull = R.measurement(n, X)
    y = np.zeros(q)
    y[0], y[1] = ox - X[0], oy - X[1]
    y[2] = float(R.normalize_angle(z2 - X[2]))
    obs = {}
    for o_i, s in enumerate(slots):
        obs[s] = o_i
    k = 3
    for s in sorted(slots):
        rng_o = float(g["range"][obs[s]])
        brg_o = float(R.normalize_angle(g["bearing"][obs[s]]))
        y[k] = rng_o - Zfull[3 + s]
        y[k + 1] = float(R.normalize_angle(brg_o - Zfull[4 + s]))
        k += 2
    Xn = X + K @ y
    Pl = Pl - K[rb:re] @ HP
    # --- assemble and compare with the oracle's unsharded step
    full = torch.zeros((n, n), dtype=torch.float64)
    full[rb:re] = torch.from_numpy(Pl)
    dist.all_reduce(full)  # result assembly for the check only (row blocks are disjoint)
    if rank == 0:
        Pfull = full.numpy()
        from util import relerr
        Pref = g["P_colmajor"].T
        assert relerr(Xn, g["X"]) < 1e-11, relerr(Xn, g["X"])
        assert relerr(Pfull, Pref) < 1e-11, relerr(Pfull, Pref)
        assert abs(a00 - g["a"][0]) < 1e-15 and abs(a10 - g["a"][1]) < 1e-15
    dist.barrier()
    dist.destroy_process_group()


def test_row_block_exchange_reproduces_unsharded_step():
    _spawn(_rows_worker)


def test_row_block_partition_properties():
    from awesomeslam_b200 import sharding
    for n in (75, 131, 8195):
        for world in (1, 2, 4, 8):
            for align in (1, 16, 256):
                cover = []
                for r in range(world):
                    b, e = sharding.row_block(n, r, world, align)
                    assert b <= e
                    cover += list(range(b, e))
                    if r == 0:
                        assert b == 0 and e >= 3
                assert cover == list(range(n))
                assert sharding.row_owner(n, world, 0, align) == sharding.row_owner(n, world, 2, align) == 0
    assert sharding.filter_range(65536, 3, 8) == (196608, 262144)
