"""Partitioning of the two shardable paths across the GPUs of one box.

Batched filters (cfg3): independent units. Rank g owns the global filter ids
[g * per_gpu, (g + 1) * per_gpu) -- weak scaling, per-filter seeds follow the GLOBAL id, no
data-path collective.

Large-map EKF (cfg4): the covariance is sharded by row blocks. One exchange per step: every rank
contributes the rows of P named by H_o that it owns and its slice of P H_o' into a [2 q][n] buffer in
which every element has exactly ONE non-zero owner; summing the buffers of all ranks (NCCL all-reduce
over NVLink on the GPUs, gloo in the CPU tests) therefore reproduces the unsharded values exactly.

This module is host logic only (no arithmetic of the filter happens here).
"""
import numpy as np


def filter_range(per_gpu, rank, world):
    """Global filter ids of `rank` (weak scaling: `per_gpu` filters on every rank)."""
    assert 0 <= rank < world
    return rank * per_gpu, (rank + 1) * per_gpu


def row_block(n, rank, world, align=1):
    """Rows [begin, end) of the n x n covariance owned by `rank`. The three pose rows always live on
    rank 0 (large_predict_kernel updates rows 0-1 and columns 0-1 from the owner of row 0). Block
    boundaries are rounded to `align` rows; the last rank takes the remainder."""
    assert 0 <= rank < world and n >= 3

    def cut(k):
        if k <= 0:
            return 0
        if k >= world:
            return n
        c = (n * k) // world
        c = (c // align) * align
        return min(max(c, 3), n)

    return cut(rank), cut(rank + 1)


def row_owner(n, world, row, align=1):
    for g in range(world):
        b, e = row_block(n, g, world, align)
        if b <= row < e:
            return g
    raise ValueError(row)


def exchange_sum(buf, group=None):
    """Sums the exchange buffers of all ranks in place. `buf` is a torch tensor (CUDA: NCCL; CPU: gloo)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf


def exchange_contribution(P_rows, row_begin, rows, hcols, hvals, n):
    """Reference (numpy) form of one rank's exchange buffer, used by the CPU tests to pin the layout the
    CUDA gather kernel writes: E[k][j] = P[rows[k]][j] if this rank owns rows[k] else 0;
    E[q + k][i] = sum_e P[i][hcols[k][e]] * hvals[k][e] for the owned rows i else 0.
    P_rows: this rank's rows [row_begin, row_begin + P_rows.shape[0]) of P."""
    q = len(rows)
    row_end = row_begin + P_rows.shape[0]
    E = np.zeros((2 * q, n))
    for k, r in enumerate(rows):
        if row_begin <= r < row_end:
            E[k] = P_rows[r - row_begin]
        acc = np.zeros(P_rows.shape[0])
        for c, v in zip(hcols[k], hvals[k]):
            acc += P_rows[:, c] * v
        E[q + k, row_begin:row_end] = acc
    return E
