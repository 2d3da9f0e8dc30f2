"""Multi-GPU forms of snapshot placement: one process per GPU, torch.distributed (NCCL on GPUs).

Two ways to split one snapshot over W ranks:

* **pod-sharded** (``snapshot_pod_sharded``): pods are independent, so each rank scores a
  contiguous range of pods against the whole node table.  The table (<= 2.8 MB at 50 k nodes)
  is replicated with one broadcast from the source rank; there is no collective on the data
  path; results stay sharded unless ``gather=True``.
* **node-sharded** (``snapshot_node_sharded``): each rank holds a contiguous range of the node
  list (contiguity keeps "first in list order wins" meaningful) and scores ALL pods against its
  range; the per-pod winner is then an arg-max reduce over ranks under (score desc, index asc):
  an all-reduce MAX over the fp64 scores followed by an all-reduce MIN over the global indices
  of the ranks that attain that maximum.  Both are exact: scores are compared as fp64 values,
  exactly as the reference compares them (scheduler.py:547), and -1.0 / -1 stay "no node".

Greedy placement is a dependent chain over pods and does not shard without one exchange per
pod; it runs on one GPU (replicas at N > 1), see DESIGN.md.

``place_fn`` is the local placement op (default: the CUDA op); the CPU tests of this module
inject a stand-in so that the sharding / collective logic runs under gloo without a GPU.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Tuple

import torch
import torch.distributed as dist

from .tables import NODE_COLUMNS

INT64_MAX = torch.iinfo(torch.int64).max


def shard_bounds(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) of ``n_items`` for ``rank`` (first ranks get the remainder)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def _default_place():
    from . import placement
    return placement.snapshot_place


def broadcast_nodes(nodes: Dict[str, torch.Tensor], src: int = 0) -> Dict[str, torch.Tensor]:
    """Replicate the node table of rank ``src`` (every rank passes tensors of the right shape)."""
    rank, world = _world()
    if world > 1:
        for k in NODE_COLUMNS:
            dist.broadcast(nodes[k], src=src)
    return nodes


def snapshot_pod_sharded(nodes: Dict[str, torch.Tensor], req_cpu_milli: torch.Tensor,
                         req_mem_bytes: torch.Tensor, strategy=0, filter_mode: int = 1,
                         place_fn: Optional[Callable] = None, src: int = 0, gather: bool = False):
    """Every rank passes the FULL pod tensors (or at least its own range filled in) and the node
    table buffers; rank ``src``'s node table is used.  Returns this rank's (lo, hi, idx, score),
    or the full (idx, score) on every rank when ``gather`` is set."""
    place_fn = place_fn or _default_place()
    rank, world = _world()
    nodes = broadcast_nodes(nodes, src)
    p = req_cpu_milli.numel()
    lo, hi = shard_bounds(p, rank, world)
    strat = strategy[lo:hi].contiguous() if isinstance(strategy, torch.Tensor) else strategy
    idx, score = place_fn(nodes, req_cpu_milli[lo:hi].contiguous(), req_mem_bytes[lo:hi].contiguous(),
                          strat, filter_mode)
    if not gather:
        return lo, hi, idx, score
    if world == 1:
        return idx, score
    # all_gather wants equal sizes: pad every shard to the largest one, then cut the padding out
    spans = [shard_bounds(p, r, world) for r in range(world)]
    width = max(b - a for a, b in spans)
    pad_idx = torch.full((width,), -1, dtype=torch.int64, device=idx.device)
    pad_score = torch.full((width,), -1.0, dtype=torch.float64, device=idx.device)
    pad_idx[:hi - lo] = idx
    pad_score[:hi - lo] = score
    all_idx = torch.empty(world * width, dtype=torch.int64, device=idx.device)
    all_score = torch.empty(world * width, dtype=torch.float64, device=idx.device)
    dist.all_gather_into_tensor(all_idx, pad_idx)
    dist.all_gather_into_tensor(all_score, pad_score)
    full_idx = torch.cat([all_idx[r * width:r * width + (b - a)] for r, (a, b) in enumerate(spans)])
    full_score = torch.cat([all_score[r * width:r * width + (b - a)] for r, (a, b) in enumerate(spans)])
    return full_idx, full_score


def snapshot_node_sharded(local_nodes: Dict[str, torch.Tensor], node_offset: int,
                          req_cpu_milli: torch.Tensor, req_mem_bytes: torch.Tensor, strategy=0,
                          filter_mode: int = 1, place_fn: Optional[Callable] = None):
    """``local_nodes`` = this rank's contiguous slice of the node list starting at global index
    ``node_offset``; every rank passes the same pods.  Returns the global (idx, score) on every
    rank."""
    place_fn = place_fn or _default_place()
    rank, world = _world()
    idx, score = place_fn(local_nodes, req_cpu_milli, req_mem_bytes, strategy, filter_mode)
    gidx = torch.where(idx >= 0, idx + node_offset, torch.full_like(idx, INT64_MAX))
    if world == 1:
        return torch.where(gidx == INT64_MAX, torch.full_like(gidx, -1), gidx), score
    best = score.clone()
    dist.all_reduce(best, op=dist.ReduceOp.MAX)                    # per-pod best score over ranks
    cand = torch.where((score == best) & (idx >= 0), gidx, torch.full_like(gidx, INT64_MAX))
    dist.all_reduce(cand, op=dist.ReduceOp.MIN)                    # first node in list order among the best
    out_idx = torch.where(cand == INT64_MAX, torch.full_like(cand, -1), cand)
    return out_idx, best


def slice_nodes(nodes: Dict[str, torch.Tensor], lo: int, hi: int) -> Dict[str, torch.Tensor]:
    return {k: nodes[k][lo:hi].contiguous() for k in NODE_COLUMNS}
