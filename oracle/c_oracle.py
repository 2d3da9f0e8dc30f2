"""ctypes front-end of oracle/c_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

``build()`` compiles ``libkp_oracle.so`` next to this file with gcc (oracle/Makefile);
``snapshot`` / ``greedy`` take numpy int64/float64 arrays in the wire-table layout and
return numpy results.  ``snapshot`` can split pods over threads (ctypes drops the GIL),
which bench.py's CPU legs use to occupy all host cores.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "libkp_oracle.so"
_lib = None

_I64P = ctypes.POINTER(ctypes.c_int64)
_F64P = ctypes.POINTER(ctypes.c_double)
_U8P = ctypes.POINTER(ctypes.c_uint8)


def build(force: bool = False) -> Path:
    src = _HERE / "c_oracle.c"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "-B", "libkp_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(str(_LIB_PATH))
        L.kpo_snapshot_range.restype = ctypes.c_int
        L.kpo_snapshot_range.argtypes = [
            ctypes.c_int64, _I64P, _I64P, _I64P, _I64P, _U8P, _F64P, _F64P,
            ctypes.c_int64, ctypes.c_int64, _I64P, _I64P, _I64P,
            ctypes.c_int, ctypes.c_int, _I64P, _F64P]
        L.kpo_greedy.restype = ctypes.c_int
        L.kpo_greedy.argtypes = [
            ctypes.c_int64, _I64P, _I64P, _I64P, _I64P, _U8P, _F64P, _F64P,
            ctypes.c_int64, _I64P, _I64P, _I64P, ctypes.c_int, ctypes.c_int, _I64P, _F64P]
        L.kpo_score.restype = ctypes.c_double
        L.kpo_score.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                ctypes.c_int64, ctypes.c_int64]
        L.kpo_usage.restype = ctypes.c_double
        L.kpo_usage.argtypes = [ctypes.c_int64, ctypes.c_int64]
        _lib = L
    return _lib


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a, t):
    return a.ctypes.data_as(t)


def _node_args(nodes):
    """nodes: dict with the seven wire arrays (see k8s_llm_scheduler_b200.tables)."""
    arrs = (_i64(nodes["alloc_cpu_milli"]), _i64(nodes["alloc_mem_bytes"]),
            _i64(nodes["pod_count"]), _i64(nodes["max_pods"]),
            np.ascontiguousarray(nodes["ready"], dtype=np.uint8),
            _f64(nodes["cpu_usage_pct"]), _f64(nodes["mem_usage_pct"]))
    return arrs


def snapshot(nodes, req_cpu_milli, req_mem_bytes, strategy=0, strategy_ids=None,
             filter_mode=1, threads: int = 1):
    L = lib()
    a = _node_args(nodes)
    n = len(a[2])
    rc, rm = _i64(req_cpu_milli), _i64(req_mem_bytes)
    p = len(rc)
    sid = None if strategy_ids is None else _i64(strategy_ids)
    out_idx = np.empty(p, dtype=np.int64)
    out_score = np.empty(p, dtype=np.float64)

    def run(lo, hi):
        L.kpo_snapshot_range(
            n, _p(a[0], _I64P), _p(a[1], _I64P), _p(a[2], _I64P), _p(a[3], _I64P),
            _p(a[4], _U8P), _p(a[5], _F64P), _p(a[6], _F64P), lo, hi,
            _p(rc, _I64P), _p(rm, _I64P), None if sid is None else _p(sid, _I64P),
            int(strategy), int(filter_mode), _p(out_idx, _I64P), _p(out_score, _F64P))

    threads = max(1, min(int(threads), p)) if p else 1
    if threads == 1:
        run(0, p)
    else:
        cuts = np.linspace(0, p, threads + 1).astype(np.int64)
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(lambda k: run(int(cuts[k]), int(cuts[k + 1])), range(threads)))
    return out_idx, out_score


def greedy(nodes, req_cpu_milli, req_mem_bytes, strategy=0, strategy_ids=None, filter_mode=1):
    """Returns (idx, score, final_pod_count, final_cpu_usage, final_mem_usage)."""
    L = lib()
    a = _node_args(nodes)
    n = len(a[2])
    pc, cu, mu = a[2].copy(), a[5].copy(), a[6].copy()
    rc, rm = _i64(req_cpu_milli), _i64(req_mem_bytes)
    p = len(rc)
    sid = None if strategy_ids is None else _i64(strategy_ids)
    out_idx = np.empty(p, dtype=np.int64)
    out_score = np.empty(p, dtype=np.float64)
    L.kpo_greedy(n, _p(a[0], _I64P), _p(a[1], _I64P), _p(pc, _I64P), _p(a[3], _I64P),
                 _p(a[4], _U8P), _p(cu, _F64P), _p(mu, _F64P), p, _p(rc, _I64P), _p(rm, _I64P),
                 None if sid is None else _p(sid, _I64P), int(strategy), int(filter_mode),
                 _p(out_idx, _I64P), _p(out_score, _F64P))
    return out_idx, out_score, pc, cu, mu


def score(strategy: int, cu: float, mu: float, pc: int, mp: int) -> float:
    return lib().kpo_score(int(strategy), float(cu), float(mu), int(pc), int(mp))


def usage(pc: int, mp: int) -> float:
    return lib().kpo_usage(int(pc), int(mp))


def host_cores() -> int:
    return len(os.sched_getaffinity(0))
