"""Query-batch sharding across ranks and the packed (F, query id) min-reduction.

Replaces the reference's round-robin ownership (`kidx = rank; kidx += world_size`,
/root/reference/main.cu:303-307) and its Pair gather + scatter + two-pass argmin
(main.cu:324-397).  Groups are independent and the graph is replicated, so the data path
needs no collective; the only exchange is ONE 8-byte min all-reduce of
``(F << 24) | global_query_id`` — equal F resolve to the smaller id, i.e. the reference's
lowest-index tie-break — carried by ``torch.distributed`` (NCCL over NVLink on GPUs,
gloo in the CPU tests)."""
from __future__ import annotations

import numpy as np

from . import argmin_packed, shard_range, unpack_argmin

UINT64_MAX = (1 << 64) - 1


def allreduce_packed_min(packed: int, device=None) -> int:
    """Min over ranks of an unsigned 64-bit packed word.  The word is < 2^63 whenever a
    rank owns a group (F < 2^39), and UINT64_MAX (no groups) maps to INT64_MAX, so a
    signed int64 MIN — the dtype every backend reduces — orders it correctly."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return packed
    as_signed = (1 << 63) - 1 if packed == UINT64_MAX else packed
    if as_signed >= 1 << 63:
        raise OverflowError("packed argmin word does not fit 63 bits")
    t = torch.tensor([as_signed], dtype=torch.int64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    v = int(t.item())
    return UINT64_MAX if v == (1 << 63) - 1 else v


def _allreduce_min_i64(value: int, device=None) -> int:
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.int64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return int(t.item())


def distributed_argmin(local_sums: np.ndarray, first_qid: int, device=None):
    """(min_k, min_f) over all ranks' groups; (-1, -1) when there are none (K = 0).
    One packed 8-byte reduction; if a sum or an id does not fit the packed word on any rank
    (F >= 2^39 or id >= 2^24), two plain reductions instead: min F, then the smallest id
    among its holders — the same lowest-index tie-break (main.cu:391-395)."""
    from . import MsbfsError, argmin

    local_sums = np.asarray(local_sums, dtype=np.int64)
    try:
        packed, ok = argmin_packed(local_sums, first_qid), 1
    except MsbfsError:
        packed, ok = UINT64_MAX, 0
    if _allreduce_min_i64(ok, device) == 1:
        return unpack_argmin(allreduce_packed_min(packed, device))
    big = (1 << 63) - 1
    lk, lf = argmin(local_sums)
    gf = _allreduce_min_i64(lf if lk >= 0 else big, device)
    gk = _allreduce_min_i64(first_qid + lk if (lk >= 0 and lf == gf) else big, device)
    return (-1, -1) if gk == big else (gk, gf)


__all__ = ["shard_range", "allreduce_packed_min", "distributed_argmin", "UINT64_MAX"]
