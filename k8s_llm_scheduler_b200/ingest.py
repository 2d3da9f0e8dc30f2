"""Table ingest on the GPU: Kubernetes quantity strings -> placement table columns.

Host side of ``kp_ingest_pod_requests`` / ``kp_ingest_nodes`` (include/k8s_place_b200.h).
The reference does this work in ``_parse_cpu`` / ``_parse_memory`` (scheduler.py:172-187),
``get_node_metrics`` (:139-151) and ``_convert_pod_to_spec`` (:731-764), one Python object at
a time.  Here the strings of a whole batch are packed into Arrow-style columns (bytes +
offsets), parsed by one kernel, and only the rows the device flags -- strings outside its
grammar, or values with no exact wire form -- are redone on the host with the literal mirror
in ``tables.py`` (real ``float()``, same exceptions as the reference).  No CPU fallback for
the bulk: without the CUDA library this module raises.
"""
from __future__ import annotations

from ctypes import byref, c_void_p
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import ROW_CPU_INEXACT, ROW_HOST, ROW_MEM_INEXACT, KpNodeStrings, KpPodRequestStrings, check
from .tables import (WIRE_LIMIT, NodeTable, PodTable, _record_walker, cores_to_milli, gib_to_bytes, is_ready, parse_cpu,
                     parse_memory, pod_requests, usage_percent)


def pack_strings(strs: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    """One uint8 buffer + int64 offsets (len + 1) for a list of strings (UTF-8)."""
    n = len(strs)
    off = np.zeros(n + 1, dtype=np.int64)
    joined = "".join(strs)
    if joined.isascii():
        if n:
            np.cumsum(np.fromiter(map(len, strs), dtype=np.int64, count=n), out=off[1:])
        data = np.frombuffer(bytearray(joined, "ascii"), dtype=np.uint8)
    else:
        enc = [s.encode("utf-8") for s in strs]
        if n:
            np.cumsum(np.fromiter(map(len, enc), dtype=np.int64, count=n), out=off[1:])
        data = np.frombuffer(bytearray(b"".join(enc)), dtype=np.uint8)
    return data, off


def _dev(a: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a)).to(device, non_blocking=False)


def _ptr(t: Optional[torch.Tensor]):
    return c_void_p(t.data_ptr()) if t is not None and t.numel() else None


@dataclass
class PodIngest:
    table: PodTable                  # wire columns for the placement kernels
    cpu_request: np.ndarray          # PodSpec.cpu_request, fp64 cores, the reference's bits
    memory_request: np.ndarray       # PodSpec.memory_request, fp64 GiB
    status: np.ndarray               # device flags per pod (before the host redo)
    host_rows: int                   # pods redone by the host mirror


@dataclass
class NodeIngest:
    table: NodeTable
    available_cpu: np.ndarray        # NodeMetrics.available_cpu
    available_memory: np.ndarray     # NodeMetrics.available_memory
    status: np.ndarray
    host_rows: int


def flatten_pod_requests(pods: Sequence[Sequence[Optional[Dict[str, Any]]]]):
    """Per-pod lists of container ``requests`` mappings -> CSR of container rows.
    Containers without requests are dropped (scheduler.py:738); a missing key reads as "0"
    (:739-740).  Pods holding a non-string quantity are listed in ``odd`` for the host."""
    begin = np.zeros(len(pods) + 1, dtype=np.int64)
    cpu: List[str] = []
    mem: List[str] = []
    odd: List[int] = []
    for j, containers in enumerate(pods):
        for req in containers:
            if not req:
                continue
            c, m = req.get("cpu", "0"), req.get("memory", "0")
            if not isinstance(c, str) or not isinstance(m, str):
                odd.append(j)
                c, m = "0", "0"
            cpu.append(c)
            mem.append(m)
        begin[j + 1] = len(cpu)
    return begin, cpu, mem, odd


def _round_inexact(status, flag, values, scale, wire, scalar_fn, exact) -> None:
    """Rows the device parsed (fp64 value right) but flagged as having no exact wire value:
    ``exact`` -> the scalar converter raises for the first one, as ``tables`` does; otherwise
    round half-to-even like ``int(round(x))`` there, vectorised."""
    rows = np.nonzero(((status & flag) != 0) & ((status & ROW_HOST) == 0))[0]
    if rows.size == 0:
        return
    if exact:                           # raises at the first truly inexact row (a -0.0 sum passes as 0)
        for j in rows.tolist():
            wire[j] = scalar_fn(float(values[j]), True)
        return
    scaled = values[rows] * scale
    off_range = ~(np.abs(scaled) < WIRE_LIMIT)
    if off_range.any():
        scalar_fn(float(values[rows[np.argmax(off_range)]]), False)      # raises ValueError
    wire[rows] = np.rint(scaled).astype(np.int64)


def pack_pod_requests(pods: Sequence[Sequence[Optional[Dict[str, Any]]]]):
    """``flatten_pod_requests`` + ``pack_strings`` -> (begin, cpu_text, cpu_off, mem_text, mem_off, odd).
    One pass in C when the record walker is built (csrc/kp_pyobj.c), else the Python functions above."""
    walker = _record_walker()
    if walker is not None and type(pods) in (list, tuple):
        b, ct, co, mt, mo, odd = walker.pack_pod_requests(pods)
        u8 = lambda x: np.frombuffer(x, dtype=np.uint8)      # noqa: E731
        i64 = lambda x: np.frombuffer(x, dtype=np.int64)     # noqa: E731
        return i64(b), u8(ct), i64(co), u8(mt), i64(mo), odd
    begin, cpu, mem, odd = flatten_pod_requests(pods)
    cpu_text, cpu_off = pack_strings(cpu)
    mem_text, mem_off = pack_strings(mem)
    return begin, cpu_text, cpu_off, mem_text, mem_off, odd


def ingest_pod_requests(pods: Sequence[Sequence[Optional[Dict[str, Any]]]], device="cuda:0",
                        exact: bool = True, strategy_ids=None, redo: bool = True) -> PodIngest:
    """GPU form of ``PodTable.from_container_requests`` (same inputs, same table, same errors).
    ``redo=False`` returns the raw device results (flagged rows left as the device wrote them)."""
    lib = _lib.load()
    n = len(pods)
    begin, cpu_text, cpu_off, mem_text, mem_off, odd = pack_pod_requests(pods)
    dev = torch.device(device)
    d = [_dev(a, dev) for a in (begin, cpu_text, cpu_off, mem_text, mem_off)]
    out_milli = torch.empty(n, dtype=torch.int64, device=dev)
    out_bytes = torch.empty(n, dtype=torch.int64, device=dev)
    out_cpu = torch.empty(n, dtype=torch.float64, device=dev)
    out_mem = torch.empty(n, dtype=torch.float64, device=dev)
    out_status = torch.empty(n, dtype=torch.uint8, device=dev)
    st = KpPodRequestStrings(n, d[0].data_ptr(), d[1].data_ptr() if d[1].numel() else None, d[2].data_ptr(),
                             d[3].data_ptr() if d[3].numel() else None, d[4].data_ptr())
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        check(lib.kp_ingest_pod_requests(byref(st), _ptr(out_milli), _ptr(out_bytes), _ptr(out_cpu),
                                         _ptr(out_mem), _ptr(out_status), c_void_p(stream)),
              "kp_ingest_pod_requests")
    milli, nbytes = out_milli.cpu().numpy(), out_bytes.cpu().numpy()
    cpu_req, mem_req, status = out_cpu.cpu().numpy(), out_mem.cpu().numpy(), out_status.cpu().numpy()
    if odd:
        status = status.copy()
        status[np.asarray(odd, dtype=np.int64)] |= ROW_HOST
    host = np.nonzero(status & ROW_HOST)[0] if redo else np.empty(0, np.int64)
    for j in host.tolist():                  # literal mirror: real float(), the reference's exceptions
        c, m = pod_requests(pods[j])
        cpu_req[j], mem_req[j] = c, m
        milli[j], nbytes[j] = cores_to_milli(c, exact), gib_to_bytes(m, exact)
    if redo:                                 # values parsed on the device that have no exact wire form
        _round_inexact(status, ROW_CPU_INEXACT, cpu_req, 1000.0, milli, cores_to_milli, exact)
        _round_inexact(status, ROW_MEM_INEXACT, mem_req, float(2 ** 30), nbytes, gib_to_bytes, exact)
    redo = host
    sid = None if strategy_ids is None else np.ascontiguousarray(strategy_ids, dtype=np.int64)
    table = PodTable(milli, nbytes, sid, [f"pod-{j}" for j in range(n)])
    return PodIngest(table, cpu_req, mem_req, status, len(redo))


def ingest_nodes(rows: Sequence[Dict[str, Any]], device="cuda:0", exact: bool = True,
                 redo: bool = True) -> NodeIngest:
    """GPU form of ``NodeTable.from_allocatable``: rows {name, cpu, memory, pods, pod_count,
    conditions} with the allocatable quantities as the API's strings."""
    lib = _lib.load()
    n = len(rows)
    cols = []
    odd = []
    for key in ("cpu", "memory", "pods"):
        vals = []
        for i, r in enumerate(rows):
            v = r[key]
            if not isinstance(v, str):
                odd.append(i)
                v = "0"
            vals.append(v)
        cols.append(pack_strings(vals))
    pod_count = np.fromiter((int(r["pod_count"]) for r in rows), dtype=np.int64, count=n)
    ready = np.fromiter((1 if is_ready(r.get("conditions", [{"type": "Ready", "status": "True"}])) else 0
                         for r in rows), dtype=np.uint8, count=n)
    dev = torch.device(device)
    d = [_dev(a, dev) for pair in cols for a in pair] + [_dev(pod_count, dev)]
    i64 = lambda: torch.empty(n, dtype=torch.int64, device=dev)      # noqa: E731
    f64 = lambda: torch.empty(n, dtype=torch.float64, device=dev)    # noqa: E731
    o_milli, o_bytes, o_mp, o_cu, o_mu, o_ac, o_am = i64(), i64(), i64(), f64(), f64(), f64(), f64()
    o_status = torch.empty(n, dtype=torch.uint8, device=dev)
    st = KpNodeStrings(n, *[t.data_ptr() if t.numel() else None for t in d])
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        check(lib.kp_ingest_nodes(byref(st), _ptr(o_milli), _ptr(o_bytes), _ptr(o_mp), _ptr(o_cu), _ptr(o_mu),
                                  _ptr(o_ac), _ptr(o_am), _ptr(o_status), c_void_p(stream)), "kp_ingest_nodes")
    milli, nbytes, mp = o_milli.cpu().numpy(), o_bytes.cpu().numpy(), o_mp.cpu().numpy()
    cu, mu, ac, am = o_cu.cpu().numpy(), o_mu.cpu().numpy(), o_ac.cpu().numpy(), o_am.cpu().numpy()
    status = o_status.cpu().numpy()
    if odd:
        status = status.copy()
        status[np.asarray(odd, dtype=np.int64)] |= ROW_HOST
    redo = set(np.nonzero(status)[0].tolist()) if redo else set()
    for i in sorted(redo):
        r = rows[i]
        max_pods = int(r["pods"])
        if not abs(max_pods) < WIRE_LIMIT:
            raise ValueError(f"allocatable pods {r['pods']!r} is outside the wire format's range (|value| < 2^53)")
        mp[i] = max_pods
        cu[i] = mu[i] = usage_percent(int(pod_count[i]), int(mp[i]))
        ac[i], am[i] = parse_cpu(r["cpu"]), parse_memory(r["memory"])
        milli[i], nbytes[i] = cores_to_milli(ac[i], exact), gib_to_bytes(am[i], exact)
    table = NodeTable(milli, nbytes, pod_count, mp, ready, cu, mu, [r.get("name") for r in rows])
    return NodeIngest(table, ac, am, status, len(redo))


__all__ = ["pack_strings", "flatten_pod_requests", "ingest_pod_requests", "ingest_nodes", "PodIngest",
           "NodeIngest", "ROW_HOST"]
