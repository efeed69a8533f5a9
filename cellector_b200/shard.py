"""Host-side sharding of cells across ranks: contiguous cell ranges balanced by nonzeros, the
rank-local COO a rank hands to clx_load_coo, and the gathers that put per-cell results back in global order.
Collectives go through torch.distributed, so the same code runs over NCCL on GPUs and over gloo in CPU tests."""
from __future__ import annotations

import numpy as np


def partition_cells(row_nnz: np.ndarray, world: int) -> np.ndarray:
    """Boundaries b[0..world] of contiguous cell ranges with about equal nonzeros: rank r owns [b[r], b[r+1])."""
    row_nnz = np.asarray(row_nnz, np.int64)
    n = row_nnz.shape[0]
    csum = np.concatenate([[0], np.cumsum(row_nnz)])
    total = int(csum[-1])
    bounds = [0]
    for r in range(1, world):
        target = total * r // world
        b = int(np.searchsorted(csum, target, side="left"))
        bounds.append(min(max(b, bounds[-1]), n))
    bounds.append(n)
    return np.asarray(bounds, np.int64)


def shard_coo(locus, cell, ref, alt, bounds, rank):
    """Entries of the cells of `rank`, in file order, with rank-local cell ids."""
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    cell = np.asarray(cell)
    keep = (cell >= lo) & (cell < hi)
    return (np.ascontiguousarray(np.asarray(locus)[keep], np.uint32), np.ascontiguousarray(cell[keep] - lo, np.uint32),
            np.ascontiguousarray(np.asarray(ref)[keep], np.uint32), np.ascontiguousarray(np.asarray(alt)[keep], np.uint32), hi - lo)


def allreduce_sum(a: np.ndarray) -> np.ndarray:
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.ascontiguousarray(a).astype(np.int64))
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t)
    return t.cpu().numpy()


def gather_cells(local: np.ndarray, bounds) -> np.ndarray:
    """Concatenate per-rank per-cell vectors into global cell order (ranks own contiguous ranges)."""
    import torch.distributed as dist
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, np.asarray(local))
    for r, p in enumerate(parts):
        assert p.shape[0] == int(bounds[r + 1] - bounds[r]), "rank result does not match its cell range"
    return np.concatenate(parts)


def locus_filter_counts(locus, ref, alt, n_loci0):
    """Per locus: entries with ref > 0 and with alt > 0 of THIS shard (load_data.rs:268-269); summed over ranks and
    compared with min_ref / min_alt they give the loci every rank keeps."""
    locus = np.asarray(locus, np.int64)
    cr = np.bincount(locus[np.asarray(ref) > 0], minlength=n_loci0)
    ca = np.bincount(locus[np.asarray(alt) > 0], minlength=n_loci0)
    return np.stack([cr, ca], axis=1)
