"""Placement ops over CUDA tensors, and the host-buffer session.

``snapshot_place`` / ``greedy_place`` / ``node_scores`` take int64 / float64 CUDA tensors and launch on
torch's current stream.  They are registered PyTorch operators -- ``torch.ops.k8s_place.*`` (csrc/kp_torch_ext.cpp,
``TORCH_LIBRARY(k8s_place, ...)``, CUDA dispatch key only), the extension form the north star names -- which call
the C ABI of libk8s_place_b200.so; the functions here are their Python signatures (a dict of node columns instead
of seven positional tensors).  ``Session`` takes host numpy arrays (or pinned torch tensors) through ctypes and is
what the reference-facing adapter uses.  PyTorch is plumbing here: device memory, streams, the dispatcher.
Nothing in this module computes a placement on the CPU.
"""
from __future__ import annotations

import ctypes
from ctypes import byref, c_int64, c_void_p
from typing import Dict, Optional, Tuple, Union

import numpy as np
import torch

from . import _lib
from ._lib import FILTER_CAPACITY, FILTER_READY, KpNodeTable, KpPodTable, check
from .tables import NODE_COLUMNS, NodeTable, PodTable

_NODE_DTYPES = {"alloc_cpu_milli": torch.int64, "alloc_mem_bytes": torch.int64,
                "pod_count": torch.int64, "max_pods": torch.int64, "ready": torch.uint8,
                "cpu_usage_pct": torch.float64, "mem_usage_pct": torch.float64}


_OPS = [None]


def ops():
    """``torch.ops.k8s_place`` with the extension library loaded (raises if it has not been built: no fallback)."""
    if _OPS[0] is None:
        from .build import TORCH_EXT_PATH
        _lib.load()                                   # the C-ABI library the extension links against
        if not TORCH_EXT_PATH.exists():
            raise ImportError(f"{TORCH_EXT_PATH} not found: build it with `python -m k8s_llm_scheduler_b200.build`")
        torch.ops.load_library(str(TORCH_EXT_PATH))
        _OPS[0] = torch.ops.k8s_place
    return _OPS[0]


def _cols(nodes: Dict[str, torch.Tensor]):
    return tuple(nodes[k] for k in NODE_COLUMNS)


def _dev_col(t: torch.Tensor, name: str, dtype: torch.dtype, device: torch.device, n: int):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name}: expected a CUDA tensor")
    if t.device != device:
        raise ValueError(f"{name}: on {t.device}, expected {device}")
    if t.dtype != dtype:
        raise TypeError(f"{name}: dtype {t.dtype}, expected {dtype}")
    if t.dim() != 1 or t.numel() != n:
        raise ValueError(f"{name}: shape {tuple(t.shape)}, expected ({n},)")
    if not t.is_contiguous():
        raise ValueError(f"{name}: must be contiguous")
    return t.data_ptr()


def nodes_to_device(table: Union[NodeTable, Dict[str, np.ndarray]], device) -> Dict[str, torch.Tensor]:
    cols = table.columns() if isinstance(table, NodeTable) else table
    dev = torch.device(device)
    return {k: torch.as_tensor(np.ascontiguousarray(cols[k])).to(_NODE_DTYPES[k]).to(dev)
            for k in NODE_COLUMNS}


def _node_struct(nodes: Dict[str, torch.Tensor]) -> Tuple[KpNodeTable, torch.device, int]:
    pc = nodes["pod_count"]
    device, n = pc.device, pc.numel()
    st = KpNodeTable()
    st.n_nodes = n
    for k in NODE_COLUMNS:
        setattr(st, k, _dev_col(nodes[k], k, _NODE_DTYPES[k], device, n))
    return st, device, n


def _pod_struct(req_cpu_milli, req_mem_bytes, strategy, device):
    p = req_cpu_milli.numel()
    st = KpPodTable()
    st.n_pods = p
    st.req_cpu_milli = _dev_col(req_cpu_milli, "req_cpu_milli", torch.int64, device, p)
    st.req_mem_bytes = _dev_col(req_mem_bytes, "req_mem_bytes", torch.int64, device, p)
    if isinstance(strategy, torch.Tensor):
        st.strategy_ids = _dev_col(strategy, "strategy_ids", torch.int64, device, p)
        return st, p, 0, True
    st.strategy_ids = None
    return st, p, int(strategy), False


def _workspace(nbytes: int, device, workspace: Optional[torch.Tensor]) -> torch.Tensor:
    if workspace is None:
        return torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
    if workspace.device != device or workspace.dtype != torch.uint8 or workspace.numel() < nbytes:
        raise ValueError(f"workspace must be a uint8 tensor on {device} with >= {nbytes} bytes")
    return workspace


def _strategy_args(strategy):
    if isinstance(strategy, torch.Tensor):
        return strategy, 0
    return None, int(strategy)


def snapshot_workspace_bytes(n_nodes: int, n_pods: int, per_pod_strategy: bool) -> int:
    return int(_lib.load().kp_snapshot_workspace_bytes(n_nodes, n_pods, int(per_pod_strategy)))


def greedy_workspace_bytes(n_nodes: int, n_pods: int) -> int:
    return int(_lib.load().kp_greedy_workspace_bytes(n_nodes, n_pods))


def snapshot_place(nodes: Dict[str, torch.Tensor], req_cpu_milli: torch.Tensor,
                   req_mem_bytes: torch.Tensor, strategy: Union[int, torch.Tensor] = 0,
                   filter_mode: int = FILTER_CAPACITY, workspace: Optional[torch.Tensor] = None,
                   out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """P independent fallback decisions (scheduler.py:521-559) against one node table.

    Returns (node_idx int64[P], score float64[P]); idx -1 / score -1.0 = no node.
    """
    sid, strat = _strategy_args(strategy)
    op = ops()
    if out is None and workspace is None:
        return op.snapshot_place(*_cols(nodes), req_cpu_milli, req_mem_bytes, sid, strat, int(filter_mode))
    device, n, p = req_cpu_milli.device, nodes["pod_count"].numel(), req_cpu_milli.numel()
    if out is None:
        out = (torch.empty(p, dtype=torch.int64, device=device), torch.empty(p, dtype=torch.float64, device=device))
    ws = _workspace(snapshot_workspace_bytes(n, p, sid is not None), device, workspace)
    op.snapshot_place_out(*_cols(nodes), req_cpu_milli, req_mem_bytes, sid, strat, int(filter_mode), out[0], out[1], ws)
    return out[0], out[1]


def snapshot_place_again(nodes: Dict[str, torch.Tensor], req_cpu_milli: torch.Tensor, req_mem_bytes: torch.Tensor,
                         strategy: int, workspace: torch.Tensor, out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """More pods against the SAME node table: ``workspace`` still holds the node plan of an earlier
    ``snapshot_place(nodes, ..., strategy, FILTER_CAPACITY, workspace=workspace)`` call, so only the placement kernel
    runs (kp_snapshot_place_again; the caller vouches for "same table, same workspace")."""
    lib = _lib.load()
    nst, device, n = _node_struct(nodes)
    pst, p, strat, per_pod = _pod_struct(req_cpu_milli, req_mem_bytes, int(strategy), device)
    if out is None:
        out = (torch.empty(p, dtype=torch.int64, device=device), torch.empty(p, dtype=torch.float64, device=device))
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device).cuda_stream
        check(lib.kp_snapshot_place_again(byref(nst), byref(pst), strat, FILTER_CAPACITY, out[0].data_ptr(), out[1].data_ptr(),
                                          workspace.data_ptr(), workspace.numel(), c_void_p(stream)), "kp_snapshot_place_again")
    return out[0], out[1]


def snapshot_place_pruned(nodes: Dict[str, torch.Tensor], req_cpu_milli: torch.Tensor, req_mem_bytes: torch.Tensor,
                          strategy: Union[int, torch.Tensor] = 0, filter_mode: int = FILTER_CAPACITY, probe_depth: int = 32,
                          workspace: Optional[torch.Tensor] = None, out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
                          scanned: bool = False):
    """``snapshot_place`` with the first-feasible probe (kp_snapshot_place_pruned): the same (node_idx, score), but a
    pod that fits one of the first ``probe_depth`` best-first nodes of its strategy is answered without a table scan.
    ``scanned=True`` also returns how many pods per strategy still took the full scan (synchronises)."""
    if workspace is None and out is None and not scanned:        # the registered operator (torch.ops.k8s_place.snapshot_place_pruned)
        sid, strat = _strategy_args(strategy)
        return ops().snapshot_place_pruned(*_cols(nodes), req_cpu_milli, req_mem_bytes, sid, strat, int(filter_mode), int(probe_depth))
    lib = _lib.load()
    nst, device, n = _node_struct(nodes)
    pst, p, strat, per_pod = _pod_struct(req_cpu_milli, req_mem_bytes, strategy, device)
    ws = _workspace(snapshot_workspace_bytes(n, p, True), device, workspace)
    if out is None:
        out = (torch.empty(p, dtype=torch.int64, device=device), torch.empty(p, dtype=torch.float64, device=device))
    with torch.cuda.device(device):
        stream = c_void_p(torch.cuda.current_stream(device).cuda_stream)
        check(lib.kp_snapshot_place_pruned(byref(nst), byref(pst), strat, int(filter_mode), int(probe_depth), out[0].data_ptr(),
                                           out[1].data_ptr(), ws.data_ptr(), ws.numel(), stream), "kp_snapshot_place_pruned")
        if scanned:
            cnt = (c_int64 * 3)()
            check(lib.kp_snapshot_scanned_pods(ws.data_ptr(), stream, cnt), "kp_snapshot_scanned_pods")
            return out[0], out[1], [int(c) for c in cnt]
    return out[0], out[1]


def greedy_place(nodes: Dict[str, torch.Tensor], req_cpu_milli: torch.Tensor,
                 req_mem_bytes: torch.Tensor, strategy: Union[int, torch.Tensor] = 0,
                 filter_mode: int = FILTER_CAPACITY, workspace: Optional[torch.Tensor] = None):
    """Pods in arrival order with commit (scheduler.py:690-729, :144-151).

    Returns (node_idx, score, final_pod_count, final_cpu_usage_pct, final_mem_usage_pct).
    The input node tensors are not modified.
    """
    sid, strat = _strategy_args(strategy)
    return ops().greedy_place(*_cols(nodes), req_cpu_milli, req_mem_bytes, sid, strat, int(filter_mode), workspace)


def node_scores(nodes: Dict[str, torch.Tensor], strategy: int = 0, filter_mode: int = FILTER_CAPACITY):
    """Per-node score, free cpu cores, free memory GiB, eligibility (scheduler.py:537-545, :232-233)."""
    return ops().node_scores(*_cols(nodes), int(strategy), int(filter_mode))


def launch_count() -> int:
    return int(_lib.load().kp_launch_count())


# ---------------------------------------------------------------------------------------
# host-buffer session
# ---------------------------------------------------------------------------------------
def _host_ptr(a, dtype) -> Tuple[int, object]:
    """Pointer to host data of a numpy array or CPU torch tensor (kept alive by the caller)."""
    if isinstance(a, torch.Tensor):
        if a.is_cuda:
            raise TypeError("Session takes host memory; use snapshot_place/greedy_place for CUDA tensors")
        want = {np.int64: torch.int64, np.float64: torch.float64, np.uint8: torch.uint8}[dtype]
        if a.dtype != want or not a.is_contiguous():
            a = a.to(want).contiguous()
        return a.data_ptr(), a
    arr = np.ascontiguousarray(a, dtype=dtype)
    return arr.ctypes.data, arr


class Session:
    """Owns device buffers for up to (max_nodes, max_pods); every call takes HOST arrays,
    copies them in, runs the CUDA path, copies results out and waits."""

    def __init__(self, device: int = 0, max_nodes: int = 65536, max_pods: int = 1 << 20):
        self._lib = _lib.load()
        self._h = c_void_p()
        self.device, self.max_nodes, self.max_pods = device, max_nodes, max_pods
        self._resident = -1
        self._memo = {}
        self._d_idx, self._d_score = c_int64(-1), ctypes.c_double(-1.0)      # decide(): answer cells, made once
        self._d_idx_ref, self._d_score_ref = byref(self._d_idx), byref(self._d_score)
        check(self._lib.kp_session_create(device, max_nodes, max_pods, byref(self._h)),
              "kp_session_create")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.kp_session_destroy(self._h)
            self._h = c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @staticmethod
    def _node_struct(nodes):
        cols = nodes.columns() if isinstance(nodes, NodeTable) else nodes
        keep = []
        st = KpNodeTable()
        st.n_nodes = len(cols["pod_count"])
        for k in NODE_COLUMNS:
            dt = {torch.int64: np.int64, torch.float64: np.float64, torch.uint8: np.uint8}[_NODE_DTYPES[k]]
            ptr, ref = _host_ptr(cols[k], dt)
            keep.append(ref)
            setattr(st, k, ptr)
        return st, keep

    @staticmethod
    def _pod_struct(pods):
        if isinstance(pods, PodTable):
            rc, rm, sid = pods.req_cpu_milli, pods.req_mem_bytes, pods.strategy_ids
        else:
            rc, rm, sid = pods
        keep = []
        st = KpPodTable()
        p1, r1 = _host_ptr(rc, np.int64)
        p2, r2 = _host_ptr(rm, np.int64)
        keep += [r1, r2]
        st.n_pods = len(r1)
        st.req_cpu_milli, st.req_mem_bytes = p1, p2
        st.strategy_ids = None
        if sid is not None:
            p3, r3 = _host_ptr(sid, np.int64)
            keep.append(r3)
            st.strategy_ids = p3
        return st, keep

    def set_prune_depth(self, probe_depth: int) -> None:
        """First-feasible probe for this session's snapshot() calls (kp_session_set_prune_depth; 0 = off)."""
        check(self._lib.kp_session_set_prune_depth(self._h, int(probe_depth)), "kp_session_set_prune_depth")

    def set_nodes(self, nodes) -> None:
        st, keep = self._node_struct(nodes)
        check(self._lib.kp_session_set_nodes(self._h, byref(st)), "kp_session_set_nodes")
        self._resident = st.n_nodes

    # Repeated calls with the very same torch buffers (a loop over pinned tables) skip rebuilding the
    # ctypes structs: the memo holds the objects themselves and is honoured only while every tensor is
    # the same object at the same address.
    def _memo_get(self, slot, owner, tensors):
        hit = self._memo.get(slot)
        if hit is None or hit[0] is not owner or len(hit[1]) != len(tensors):
            return None
        for t, (ref, ptr, count) in zip(tensors, hit[1]):
            if t is not ref or t.data_ptr() != ptr or t.numel() != count:
                return None
        return hit[2]

    def _memo_put(self, slot, owner, tensors, keep, value):
        if all(type(t) is torch.Tensor and t is k for t, k in zip(tensors, keep)):
            self._memo[slot] = (owner, [(t, t.data_ptr(), t.numel()) for t in tensors], value)

    # Both helpers return (struct, keep): `keep` holds whatever arrays the struct points into --
    # including temporaries made by _host_ptr when an input needed a dtype / layout conversion --
    # and the caller keeps it alive until the C call has returned.
    def _node_struct_memo(self, nodes):
        if type(nodes) is not dict:
            return self._node_struct(nodes)
        tensors = [nodes[k] for k in NODE_COLUMNS]
        hit = self._memo_get("nodes", nodes, tensors)
        if hit is None:
            hit = self._node_struct(nodes)
            self._memo_put("nodes", nodes, tensors, hit[1], hit)
        return hit

    def _pod_struct_memo(self, pods):
        if type(pods) is not tuple:
            return self._pod_struct(pods)
        tensors = [t for t in pods if t is not None]
        hit = self._memo_get("pods", pods, tensors)
        if hit is None:
            hit = self._pod_struct(pods)
            self._memo_put("pods", pods, tensors, hit[1], hit)
        return hit

    def snapshot(self, nodes, pods, strategy: int = 0, filter_mode: int = FILTER_CAPACITY, out=None):
        """nodes=None reuses the table kept by set_nodes().  Returns (idx, score) numpy arrays,
        or fills ``out`` = (idx, score) host buffers (numpy or pinned torch)."""
        nst, _keep_nodes = self._node_struct_memo(nodes) if nodes is not None else (None, None)
        pst, _keep_pods = self._pod_struct_memo(pods)
        p = pst.n_pods
        if out is None:
            out = (np.empty(p, np.int64), np.empty(p, np.float64))
        ptrs = self._memo_get("out", out, list(out)) if type(out) is tuple else None
        if ptrs is None:
            ip, iref = _host_ptr(out[0], np.int64)
            sp, sref = _host_ptr(out[1], np.float64)
            if iref is not out[0] or sref is not out[1]:
                raise TypeError("out buffers must be contiguous int64 / float64 host arrays")
            ptrs = (ip, sp)
            if type(out) is tuple:
                self._memo_put("out", out, list(out), [iref, sref], ptrs)
        check(self._lib.kp_session_snapshot(self._h, byref(nst) if nst is not None else None,
                                            byref(pst), int(strategy), int(filter_mode), ptrs[0], ptrs[1]),
              "kp_session_snapshot")
        if nst is not None:
            self._resident = nst.n_nodes
        return out

    def decide(self, nodes, strategy: int = 0, filter_mode: int = FILTER_READY, req_cpu_milli: int = 0,
               req_mem_bytes: int = 0) -> Tuple[int, float]:
        """ONE decision for ONE pod (kp_session_decide; the reference's call shape, scheduler.py:521-559):
        (node index or -1, score).  ``nodes``: a NodeTable / column dict, or a ready-made KpNodeTable whose
        arrays the caller keeps alive (adapter.FallbackPlacer keeps one over its column buffers)."""
        nst = nodes if type(nodes) is KpNodeTable else self._node_struct_memo(nodes)
        if type(nst) is tuple:
            nst, _keep = nst
        check(self._lib.kp_session_decide(self._h, byref(nst), strategy, filter_mode, req_cpu_milli, req_mem_bytes,
                                          self._d_idx_ref, self._d_score_ref), "kp_session_decide")
        return self._d_idx.value, self._d_score.value

    def greedy(self, nodes, pods, strategy: int = 0, filter_mode: int = FILTER_CAPACITY, out=None):
        """Returns (idx, score, final_pod_count, final_cpu_usage, final_mem_usage) numpy arrays, or fills ``out`` =
        the same five host buffers (numpy, or pinned torch tensors: device-to-host copies into pageable memory are
        staged and several times slower)."""
        nst, keep = self._node_struct(nodes) if nodes is not None else (None, [])
        pst, keep2 = self._pod_struct(pods)
        p = pst.n_pods
        n = nst.n_nodes if nst is not None else self._resident
        if out is None:
            out = (np.empty(p, np.int64), np.empty(p, np.float64), np.empty(n, np.int64), np.empty(n, np.float64), np.empty(n, np.float64))
        want = ((p, np.int64), (p, np.float64), (n, np.int64), (n, np.float64), (n, np.float64))
        ptrs = []
        for buf, (count, dt) in zip(out, want):
            ptr, ref = _host_ptr(buf, dt)
            if ref is not buf or len(ref) != count:
                raise TypeError("out buffers must be contiguous host arrays: int64[p], float64[p], int64[n], float64[n], float64[n]")
            ptrs.append(ptr)
        check(self._lib.kp_session_greedy(self._h, byref(nst) if nst is not None else None, byref(pst),
                                          int(strategy), int(filter_mode), *ptrs), "kp_session_greedy")
        if nst is not None:
            self._resident = nst.n_nodes
        return out

    def last_copy_bytes(self) -> Tuple[int, int]:
        a, b = c_int64(), c_int64()
        check(self._lib.kp_session_last_copy_bytes(self._h, byref(a), byref(b)),
              "kp_session_last_copy_bytes")
        return a.value, b.value
