"""RequestCache keys for a batch of pods.

``RequestCache._generate_key`` (scheduler.py:265-271) hashes ``f"{pod_data}_{nodes_data}"`` with
MD5, where ``nodes_data`` lists every node (sorted by name) and is rebuilt and re-hashed for each
pod.  Here the node text is formatted once per batch, the per-pod prefixes are formatted on the
host (Python's float repr is the reference's own formatting), and the digests of all pods are
computed in one kernel (``kp_cache_keys``).  Returns the same hex strings the reference uses as
dictionary keys.  The cache itself stays in the reference.
"""
from __future__ import annotations

from ctypes import c_void_p
from typing import Any, List, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import check
from .ingest import pack_strings


def nodes_text(node_metrics: Sequence[Any]) -> str:
    """scheduler.py:268-269."""
    return "_".join(f"{n.name}_{n.cpu_usage_percent:.1f}_{n.memory_usage_percent:.1f}"
                    for n in sorted(node_metrics, key=lambda x: x.name))


def pod_prefix(pod_spec: Any) -> str:
    """scheduler.py:267 plus the separator of :270."""
    return f"{pod_spec.cpu_request}_{pod_spec.memory_request}_{pod_spec.priority}_"


def cache_keys(pod_specs: Sequence[Any], node_metrics: Sequence[Any], device="cuda:0") -> List[str]:
    lib = _lib.load()
    n = len(pod_specs)
    if n == 0:
        return []
    prefix, off = pack_strings([pod_prefix(p) for p in pod_specs])
    text = np.frombuffer(bytearray(nodes_text(node_metrics), "utf-8"), dtype=np.uint8)
    dev = torch.device(device)
    d_prefix, d_off = torch.from_numpy(prefix).to(dev), torch.from_numpy(off).to(dev)
    d_text = torch.from_numpy(text).to(dev) if text.size else torch.zeros(4, dtype=torch.uint8, device=dev)
    out = torch.empty((n, 16), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        check(lib.kp_cache_keys(d_prefix.data_ptr() if d_prefix.numel() else None, d_off.data_ptr(), n,
                                d_text.data_ptr(), int(text.size), out.data_ptr(), c_void_p(stream)), "kp_cache_keys")
    return [row.tobytes().hex() for row in out.cpu().numpy()]


__all__ = ["cache_keys", "nodes_text", "pod_prefix"]
