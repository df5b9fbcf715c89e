"""Multi-GPU plumbing of the hot path: one process per GPU, `torch.distributed` for the control
messages only — the probe itself never crosses devices.

* flat IBF: rank r of R owns the 64-bin column block `column_shard(bin_words, r, R)` of EVERY row
  (`trx_ibf_create(..., shard_index=r, shard_count=R)`), so row indices are rank independent;
* HIBF (small): replicated, the k-mer batch / query list is split with `unit_slice`;
* collector pruning needs one flag per item that is global over all bins: `any_alive` (all-reduce MAX of bytes);
* results: `gather_words` concatenates the per-rank column blocks of the final bin vector, `gather_hits` brings
  verify hits to rank 0 in the reference's (bin, record, start) print order.

Backend is whatever the process group was initialised with: "nccl" on the GPU box (tensors staged on the rank's
device), "gloo" in the CPU tests (tests/test_dist_gloo.py, world_size 2).
"""
from __future__ import annotations

import os

import numpy as np


def rank_world():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))


def unit_slice(n_units: int, rank: int, world: int) -> slice:
    """Contiguous, balanced split of n_units (k-mers, queries, bins) — the first n % world ranks get one more."""
    base, extra = divmod(n_units, world)
    begin = rank * base + min(rank, extra)
    return slice(begin, begin + base + (1 if rank < extra else 0))


def column_shard(bin_words: int, rank: int, world: int) -> tuple[int, int]:
    """(first word, word count) of rank's 64-bin column block; mirrors trx_ibf_create's rule (world must divide)."""
    if world < 1 or bin_words % world:
        raise ValueError(f"{world} shards do not divide {bin_words} words per row")
    per = bin_words // world
    return rank * per, per


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _device():
    import torch
    dist = _dist()
    if dist is not None and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def barrier():
    dist = _dist()
    if dist is not None:
        dist.barrier()


def max_over_ranks(x: float) -> float:
    import torch
    dist = _dist()
    if dist is None:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device=_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x: float) -> float:
    import torch
    dist = _dist()
    if dist is None:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device=_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def any_alive(alive: np.ndarray) -> np.ndarray:
    """`path_.none()` over ALL bins when the bins are column-sharded: item i lives if it lives on any rank."""
    import torch
    dist = _dist()
    a = np.ascontiguousarray(alive, dtype=np.uint8)
    if dist is None or a.size == 0:
        return a
    t = torch.from_numpy(a.copy()).to(_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.cpu().numpy()


def gather_words(local_words: np.ndarray) -> np.ndarray:
    """Final candidate-bin vector: every rank contributes its column block, all ranks get the whole vector."""
    import torch
    dist = _dist()
    w = np.ascontiguousarray(local_words, dtype=np.uint64)
    if dist is None:
        return w
    t = torch.from_numpy(w.view(np.int64).copy()).to(_device())
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return np.concatenate([o.cpu().numpy().view(np.uint64) for o in out])


def gather_hits(hits: np.ndarray):
    """hits: (n, 4) uint32 rows {bin, record, start, end} of this rank's bins.  Rank 0 gets all rows sorted by
    (bin, record, start) — the order iter_disk_search prints with -t 1; other ranks get None."""
    import torch
    dist = _dist()
    h = np.ascontiguousarray(hits, dtype=np.uint32).reshape(-1, 4)
    if dist is None:
        parts = [h]
        rank = 0
    else:
        rank, world = dist.get_rank(), dist.get_world_size()
        dev = _device()
        n = torch.tensor([h.shape[0]], dtype=torch.int64, device=dev)
        counts = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(counts, n)
        counts = [int(c.item()) for c in counts]
        cap = max(max(counts), 1)
        pad = np.zeros((cap, 4), dtype=np.int64)
        pad[:h.shape[0]] = h
        t = torch.from_numpy(pad).to(dev)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        parts = [o.cpu().numpy()[:c].astype(np.uint32) for o, c in zip(out, counts)]
    if rank != 0:
        return None
    allh = np.concatenate(parts) if parts else h
    order = np.lexsort((allh[:, 2], allh[:, 1], allh[:, 0]))
    return allh[order]
