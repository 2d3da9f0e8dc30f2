"""ctypes binding of include/k8s_place_b200.h.  There is NO fallback: if the CUDA library
is missing, or a call fails, this raises -- the product never routes through a CPU path."""
from __future__ import annotations

import ctypes
from ctypes import POINTER, c_char_p, c_double, c_int, c_int64, c_size_t, c_uint8, c_void_p
from pathlib import Path

from .build import LIB_PATH

KP_OK = 0
STRATEGY_RESOURCE_BALANCED, STRATEGY_LEAST_LOADED, STRATEGY_ROUND_ROBIN = 0, 1, 2
FILTER_READY, FILTER_CAPACITY = 0, 1

_I64P, _F64P, _U8P = POINTER(c_int64), POINTER(c_double), POINTER(c_uint8)


class KpNodeTable(ctypes.Structure):
    _fields_ = [("n_nodes", c_int64), ("alloc_cpu_milli", c_void_p), ("alloc_mem_bytes", c_void_p),
                ("pod_count", c_void_p), ("max_pods", c_void_p), ("ready", c_void_p),
                ("cpu_usage_pct", c_void_p), ("mem_usage_pct", c_void_p)]


class KpPodTable(ctypes.Structure):
    _fields_ = [("n_pods", c_int64), ("req_cpu_milli", c_void_p), ("req_mem_bytes", c_void_p),
                ("strategy_ids", c_void_p)]


class KpPodRequestStrings(ctypes.Structure):
    _fields_ = [("n_pods", c_int64), ("container_begin", c_void_p), ("cpu_text", c_void_p),
                ("cpu_offsets", c_void_p), ("mem_text", c_void_p), ("mem_offsets", c_void_p)]


class KpNodeStrings(ctypes.Structure):
    _fields_ = [("n_nodes", c_int64), ("cpu_text", c_void_p), ("cpu_offsets", c_void_p),
                ("mem_text", c_void_p), ("mem_offsets", c_void_p), ("pods_text", c_void_p),
                ("pods_offsets", c_void_p), ("pod_count", c_void_p)]


ROW_OK, ROW_CPU_INEXACT, ROW_MEM_INEXACT, ROW_HOST = 0, 1, 2, 4


class PlacementError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str):
        super().__init__(f"{where}: status {status} ({detail})")
        self.status = status


# every symbol include/k8s_place_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "kp_abi_version": (c_int, []),
    "kp_status_string": (c_char_p, [c_int]),
    "kp_last_error": (c_char_p, []),
    "kp_device_info": (c_int, [c_int, POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "kp_launch_count": (c_int64, []),
    "kp_profile_enable": (c_int, [c_int]),
    "kp_profile_last_ms": (c_int, [POINTER(ctypes.c_float), POINTER(ctypes.c_float)]),
    "kp_issue_peak": (c_int, [c_void_p, c_size_t, c_double, c_void_p, POINTER(c_double)]),
    "kp_snapshot_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "kp_snapshot_place": (c_int, [POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int,
                                  c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "kp_snapshot_place_again": (c_int, [POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int,
                                        c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "kp_snapshot_place_noalloc": (c_int, [POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int,
                                          c_void_p, c_void_p, c_void_p]),
    "kp_exchange_create": (c_int, [c_int, c_int, c_int, c_int64, c_int64, POINTER(c_void_p)]),
    "kp_exchange_handle": (c_int, [c_void_p, c_void_p]),
    "kp_exchange_connect": (c_int, [c_void_p, c_void_p]),
    "kp_exchange_connect_local": (c_int, [c_void_p, POINTER(c_void_p)]),
    "kp_exchange_destroy": (None, [c_void_p]),
    "kp_exchange_error": (c_int, [c_void_p, POINTER(c_int)]),
    "kp_exchange_set_prune_depth": (c_int, [c_void_p, c_int]),
    "kp_ns_keys": (c_int, [c_void_p, POINTER(KpNodeTable), c_int64, POINTER(KpPodTable), c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "kp_ns_place": (c_int, [c_void_p, POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "kp_ns_reduce": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "kp_snapshot_place_node_sharded": (c_int, [c_void_p, POINTER(KpNodeTable), c_int64, POINTER(KpPodTable), c_int, c_int,
                                               c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "kp_greedy_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "kp_greedy_place": (c_int, [POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int,
                                c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_size_t, c_void_p]),
    "kp_node_scores": (c_int, [POINTER(KpNodeTable), c_int, c_int, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p]),
    "kp_session_create": (c_int, [c_int, c_int64, c_int64, POINTER(c_void_p)]),
    "kp_session_destroy": (None, [c_void_p]),
    "kp_session_set_nodes": (c_int, [c_void_p, POINTER(KpNodeTable)]),
    "kp_session_snapshot": (c_int, [c_void_p, POINTER(KpNodeTable), POINTER(KpPodTable), c_int,
                                    c_int, c_void_p, c_void_p]),
    "kp_session_greedy": (c_int, [c_void_p, POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "kp_snapshot_place_pruned": (c_int, [POINTER(KpNodeTable), POINTER(KpPodTable), c_int, c_int, c_int, c_void_p, c_void_p,
                                         c_void_p, c_size_t, c_void_p]),
    "kp_snapshot_scanned_pods": (c_int, [c_void_p, c_void_p, POINTER(c_int64)]),
    "kp_session_set_prune_depth": (c_int, [c_void_p, c_int]),
    "kp_session_decide": (c_int, [c_void_p, POINTER(KpNodeTable), c_int, c_int, c_int64, c_int64,
                                  POINTER(c_int64), POINTER(c_double)]),
    "kp_session_last_copy_bytes": (c_int, [c_void_p, POINTER(c_int64), POINTER(c_int64)]),
    "kp_ingest_pod_requests": (c_int, [POINTER(KpPodRequestStrings), c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_void_p]),
    "kp_cache_keys": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p]),
    "kp_ingest_nodes": (c_int, [POINTER(KpNodeStrings), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_void_p, c_void_p]),
}

_lib = None


def load(path: Path | None = None) -> ctypes.CDLL:
    """Load the CUDA library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    import os
    p = Path(path) if path else Path(os.environ.get("KP_LIB_PATH", str(LIB_PATH)))
    if not p.exists():
        raise ImportError(
            f"{p} not found: build it with `python -m k8s_llm_scheduler_b200.build` "
            "(nvcc, sm_100a).  This package has no CPU fallback.")
    lib = ctypes.CDLL(str(p))
    for name, (restype, argtypes) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.kp_abi_version() != 1:
        raise ImportError(f"{p}: ABI version {lib.kp_abi_version()} != 1")
    if path is None:
        _lib = lib
    return lib


def check(status: int, where: str) -> None:
    if status != KP_OK:
        lib = load()
        raise PlacementError(status, where,
                             f"{lib.kp_status_string(status).decode()}: {lib.kp_last_error().decode()}")
