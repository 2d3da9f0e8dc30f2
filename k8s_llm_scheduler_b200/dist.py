"""Multi-GPU forms of snapshot placement: one process per GPU, torch.distributed (NCCL on GPUs).

Two ways to split one snapshot over W ranks:

* **pod-sharded** (``snapshot_pod_sharded``): pods are independent, so each rank scores a
  contiguous range of pods against the whole node table.  The table (<= 2.8 MB at 50 k nodes)
  is replicated with one broadcast from the source rank; there is no collective on the data
  path; results stay sharded unless ``gather=True``.
* **node-sharded**: each rank holds a contiguous range of the node list (contiguity keeps "first
  in list order wins" meaningful) and scores ALL pods against its range; the per-pod winner is
  then an arg-max reduce over ranks under (score desc, index asc).  On GPUs this is
  ``NodeShardedPlacer``: the library's own kernels exchange one 32-bit global rank position per
  pod and rank over peer-mapped memory (NVLink, CUDA IPC) and the owner of a pod shard decodes
  the winner -- no collective library and no host round trip on the data path
  (csrc/kp_exchange.cu).  ``snapshot_node_sharded`` is the same reduce written with two
  ``torch.distributed`` all-reduces (MAX over the fp64 scores, MIN over the global indices that
  attain it); it is what the CPU (gloo) tests of the sharding logic run, and a cross-check of
  the fused path on GPUs.  Both are exact: scores are compared exactly as the reference compares
  them (scheduler.py:547), and -1.0 / -1 stay "no node".

Greedy placement is a dependent chain over pods and does not shard without one exchange per
pod; it runs on one GPU (replicas at N > 1), see DESIGN.md.

``place_fn`` is the local placement op (default: the CUDA op); the CPU tests of this module
inject a stand-in so that the sharding / collective logic runs under gloo without a GPU.
"""
from __future__ import annotations

import ctypes
from typing import Callable, Dict, List, Optional, Tuple

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


# ---------------------------------------------------------------------------------------
# node-sharded snapshot over peer memory (kp_exchange.cu)
# ---------------------------------------------------------------------------------------
class NodeShardedPlacer:
    """One rank's end of the node-sharded snapshot: owns a ``kp_exchange`` (exchange memory on this GPU,
    mapped into every peer with CUDA IPC) and a placement workspace.

    ``NodeShardedPlacer(device, max_pods, max_local_nodes)`` inside an initialised process group (one
    process per GPU) exchanges the IPC handles with ``all_gather_object``; ``NodeShardedPlacer.local_group``
    builds all ranks of a group inside ONE process (several GPUs of one process, or several ranks emulated
    on one GPU for tests), linked without IPC.

    ``place`` returns this rank's pod shard ``(lo, hi, idx, score)``; indices are global (list order)."""

    def __init__(self, device, max_pods: int, max_local_nodes: int, rank: Optional[int] = None,
                 world: Optional[int] = None, connect: bool = True):
        from . import _lib
        self._lib = _lib.load()
        self._check = _lib.check
        r, w = _world()
        self.rank = r if rank is None else rank
        self.world = w if world is None else world
        self.device = torch.device(device)
        self.max_pods, self.max_local = max_pods, max_local_nodes
        self._h = ctypes.c_void_p()
        self._check(self._lib.kp_exchange_create(self.device.index or 0, self.rank, self.world, max_pods, max_local_nodes,
                                                 ctypes.byref(self._h)), "kp_exchange_create")
        self._ws: Optional[torch.Tensor] = None
        self._n_pods = 0
        self._prune = 0
        if connect and self.world > 1:
            buf = (ctypes.c_ubyte * 64)()
            self._check(self._lib.kp_exchange_handle(self._h, buf), "kp_exchange_handle")
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, bytes(buf))
            blob = b"".join(handles)
            self._check(self._lib.kp_exchange_connect(self._h, blob), "kp_exchange_connect")

    @classmethod
    def local_group(cls, devices, max_pods: int, max_local_nodes: int) -> List["NodeShardedPlacer"]:
        """All ranks of a group in this process: ``devices`` = one device per rank (they may repeat)."""
        world = len(devices)
        group = [cls(d, max_pods, max_local_nodes, rank=r, world=world, connect=False) for r, d in enumerate(devices)]
        arr = (ctypes.c_void_p * world)(*[g._h for g in group])
        for g in group:
            g._check(g._lib.kp_exchange_connect_local(g._h, arr), "kp_exchange_connect_local")
        return group

    def close(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self._h = None
            self._lib.kp_exchange_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 -- interpreter shutdown: modules may already be gone
            pass

    def _structs(self, local_nodes, req_cpu_milli, req_mem_bytes, strategy):
        from . import placement
        nst, device, n = placement._node_struct(local_nodes)
        pst, p, strat, per_pod = placement._pod_struct(req_cpu_milli, req_mem_bytes, strategy, device)
        need = placement.snapshot_workspace_bytes(n, p, per_pod or self._prune > 0)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 256), dtype=torch.uint8, device=device)
        return nst, pst, n, p, strat, per_pod

    def set_prune_depth(self, probe_depth: int) -> None:
        """Optional first-feasible probe for this rank's scan (kp_exchange_set_prune_depth; 0 = off, the default)."""
        self._check(self._lib.kp_exchange_set_prune_depth(self._h, int(probe_depth)), "kp_exchange_set_prune_depth")
        self._prune = int(probe_depth)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # the three steps, separately (ranks emulated on one device must be driven in lock step)
    def keys(self, local_nodes, node_offset: int, req_cpu_milli, req_mem_bytes, strategy, filter_mode: int = 1):
        nst, pst, n, p, strat, per_pod = self._structs(local_nodes, req_cpu_milli, req_mem_bytes, strategy)
        with torch.cuda.device(self.device):
            self._check(self._lib.kp_ns_keys(self._h, ctypes.byref(nst), int(node_offset), ctypes.byref(pst), strat, int(filter_mode),
                                             self._ws.data_ptr(), self._ws.numel(), self._stream()), "kp_ns_keys")

    def scan(self, local_nodes, req_cpu_milli, req_mem_bytes, strategy, filter_mode: int = 1):
        nst, pst, n, p, strat, per_pod = self._structs(local_nodes, req_cpu_milli, req_mem_bytes, strategy)
        self._n_pods = p
        with torch.cuda.device(self.device):
            self._check(self._lib.kp_ns_place(self._h, ctypes.byref(nst), ctypes.byref(pst), strat, int(filter_mode),
                                              self._ws.data_ptr(), self._ws.numel(), self._stream()), "kp_ns_place")

    def reduce(self, out=None):
        lo, hi = shard_bounds(self._n_pods, self.rank, self.world)
        if out is None:
            out = (torch.empty(hi - lo, dtype=torch.int64, device=self.device),
                   torch.empty(hi - lo, dtype=torch.float64, device=self.device))
        with torch.cuda.device(self.device):
            self._check(self._lib.kp_ns_reduce(self._h, out[0].data_ptr(), out[1].data_ptr(), self._stream()), "kp_ns_reduce")
        return lo, hi, out[0], out[1]

    def place(self, local_nodes: Dict[str, torch.Tensor], node_offset: int, req_cpu_milli: torch.Tensor,
              req_mem_bytes: torch.Tensor, strategy=0, filter_mode: int = 1, out=None):
        """Every rank: its node slice (CUDA tensors, list order), the slice's global start index, and the SAME pods."""
        self.keys(local_nodes, node_offset, req_cpu_milli, req_mem_bytes, strategy, filter_mode)
        self.scan(local_nodes, req_cpu_milli, req_mem_bytes, strategy, filter_mode)
        return self.reduce(out)

    def check(self):
        """After synchronising: raise if a kernel of this rank gave up waiting for a peer."""
        err = ctypes.c_int()
        self._check(self._lib.kp_exchange_error(self._h, ctypes.byref(err)), "kp_exchange_error")
        if err.value:
            raise RuntimeError(self._lib.kp_last_error().decode())
