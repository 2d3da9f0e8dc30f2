"""Host-side tables: the reference's records (scheduler.py:72-104) <-> the int64/float64
wire format of include/k8s_place_b200.h.

Parsing follows the reference bit-for-bit:

  * node CPU  (scheduler.py:172-176): only the ``m`` suffix is known; anything else is
    ``float(text)`` cores.
  * node memory (scheduler.py:178-187): ``Ki`` / ``Mi`` / ``Gi``, else plain bytes;
    ``K``/``M``/``G``/``Ti`` raise ``ValueError`` exactly like the reference's ``float()``.
  * pod requests (scheduler.py:737-753): summed over containers in fp64, and for memory
    ONLY ``Ki``/``Mi``/``Gi`` strings count -- plain-byte strings add nothing.
  * usage percent (scheduler.py:150-151): ``(pod_count / max_pods) * 50`` or 0.

The wire format is integer (milli-cores, bytes).  A float quantity q maps to the integer
m with ``m / 1000.0 == q`` (cores) or ``b / 2**30 == q`` (GiB); if no such integer exists
(e.g. a multi-container sum like 0.1 + 0.2) ``exact=True`` raises instead of rounding.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, Iterable, List, Optional, Sequence

import numpy as np

STRATEGY_IDS = {"resource_balanced": 0, "least_loaded": 1}


def strategy_id(name: str) -> int:
    """Config string -> id; every unknown string is round_robin (scheduler.py:544)."""
    return STRATEGY_IDS.get(name, 2)


# ---- the reference's record types, same field names and order (scheduler.py:72-104) ----
@dataclass
class NodeMetrics:
    name: str
    cpu_usage_percent: float
    memory_usage_percent: float
    available_cpu: float
    available_memory: float
    pod_count: int
    max_pods: int
    labels: Dict[str, str] = field(default_factory=dict)
    taints: List[Dict[str, str]] = field(default_factory=list)
    conditions: List[Dict[str, str]] = field(default_factory=list)


@dataclass
class PodSpec:
    name: str
    namespace: str
    cpu_request: float
    memory_request: float
    node_selector: Dict[str, str] = field(default_factory=dict)
    tolerations: List[Dict[str, str]] = field(default_factory=list)
    affinity_rules: Dict[str, Any] = field(default_factory=dict)
    priority: int = 0


@dataclass
class SchedulingDecision:
    selected_node: str
    confidence: float
    reasoning: str
    fallback_needed: bool = False


# ---- quantity parsing ---------------------------------------------------------------
def parse_cpu(text: str) -> float:
    return float(text[:-1]) / 1000 if text.endswith("m") else float(text)


def parse_memory(text: str) -> float:
    if text.endswith("Ki"):
        return float(text[:-2]) / 1024 / 1024
    if text.endswith("Mi"):
        return float(text[:-2]) / 1024
    if text.endswith("Gi"):
        return float(text[:-2])
    return float(text) / 1024 / 1024 / 1024


def pod_requests(containers: Iterable[Optional[Dict[str, Any]]]):
    cpu, mem = 0.0, 0.0
    for req in containers:
        if not req:
            continue
        c, m = req.get("cpu", "0"), req.get("memory", "0")
        cpu += float(c[:-1]) / 1000 if isinstance(c, str) and c.endswith("m") else float(c or 0)
        if isinstance(m, str):
            if m.endswith("Ki"):
                mem += float(m[:-2]) / 1024 / 1024
            elif m.endswith("Mi"):
                mem += float(m[:-2]) / 1024
            elif m.endswith("Gi"):
                mem += float(m[:-2])
    return cpu, mem


def usage_percent(pod_count: int, max_pods: int):
    return (pod_count / max_pods) * 50 if max_pods > 0 else 0


def is_ready(conditions: Sequence[Dict[str, str]]) -> bool:
    """scheduler.py:532-533."""
    return any(c.get("type") == "Ready" and c.get("status") == "True" for c in conditions)


WIRE_LIMIT = 2 ** 53          # |milli-cores|, |bytes| below this are exact in fp64 (kernel contract)


def _in_wire_range(x: float, what: str) -> None:
    if not abs(x) < WIRE_LIMIT:          # also catches nan / inf
        raise ValueError(f"{what} is outside the wire format's range (|value| < 2^53)")


def cores_to_milli(cores: float, exact: bool = True) -> int:
    _in_wire_range(cores * 1000, f"{cores!r} cores")
    m = int(round(cores * 1000))
    if exact and m / 1000.0 != cores:
        raise ValueError(f"{cores!r} cores is not an integer number of milli-cores in fp64")
    return m


def gib_to_bytes(gib: float, exact: bool = True) -> int:
    _in_wire_range(gib * 2 ** 30, f"{gib!r} GiB")
    b = int(round(gib * 2 ** 30))
    if exact and b / 1024 / 1024 / 1024 != gib:
        raise ValueError(f"{gib!r} GiB is not an integer number of bytes in fp64")
    return b


_WALKER = [False]


def _record_walker():
    """The optional C walker over reference-style records (built by build.build_pyobj); None when it
    has not been built or KP_NO_PYOBJ is set -- the Python loops below are then used, same results."""
    if _WALKER[0] is False:
        import os
        mod = None
        if not os.environ.get("KP_NO_PYOBJ"):
            try:
                from . import _kp_pyobj as mod          # noqa: PLC0415
            except ImportError:
                mod = None
        _WALKER[0] = mod
    return _WALKER[0]


# ---- wire tables --------------------------------------------------------------------
NODE_COLUMNS = ("alloc_cpu_milli", "alloc_mem_bytes", "pod_count", "max_pods", "ready",
                "cpu_usage_pct", "mem_usage_pct")


@dataclass
class NodeTable:
    """Structure-of-arrays node table in list order (= tie-break order)."""
    alloc_cpu_milli: np.ndarray
    alloc_mem_bytes: np.ndarray
    pod_count: np.ndarray
    max_pods: np.ndarray
    ready: np.ndarray
    cpu_usage_pct: np.ndarray
    mem_usage_pct: np.ndarray
    names: Optional[List[str]] = None

    def __len__(self) -> int:
        return len(self.pod_count)

    def columns(self) -> Dict[str, np.ndarray]:
        return {k: getattr(self, k) for k in NODE_COLUMNS}

    @classmethod
    def from_columns(cls, names=None, **cols) -> "NodeTable":
        def i64(a):
            return np.ascontiguousarray(a, dtype=np.int64)
        return cls(i64(cols["alloc_cpu_milli"]), i64(cols["alloc_mem_bytes"]), i64(cols["pod_count"]),
                   i64(cols["max_pods"]), np.ascontiguousarray(cols["ready"], dtype=np.uint8),
                   np.ascontiguousarray(cols["cpu_usage_pct"], dtype=np.float64),
                   np.ascontiguousarray(cols["mem_usage_pct"], dtype=np.float64),
                   list(names) if names is not None else None)

    @classmethod
    def from_node_metrics(cls, nodes: Sequence[Any], exact: bool = True, ready_only: bool = False) -> "NodeTable":
        """From a list of reference-style NodeMetrics records (duck-typed: the reference's
        own dataclass instances work).  ``ready_only``: the table is for the Ready-only filter, i.e. the
        reference's own decision (scheduler.py:531-549), which never reads available_cpu / available_memory:
        those fields are then not read either (zeros), whatever they hold."""
        n = len(nodes)
        t = cls(np.empty(n, np.int64), np.empty(n, np.int64), np.empty(n, np.int64),
                np.empty(n, np.int64), np.empty(n, np.uint8), np.empty(n, np.float64),
                np.empty(n, np.float64), [None] * n)
        walker = _record_walker()
        if walker is not None:               # same conversions and checks, in C (csrc/kp_pyobj.c)
            walker.node_metrics_to_columns(nodes, t.alloc_cpu_milli, t.alloc_mem_bytes, t.pod_count, t.max_pods,
                                           t.ready, t.cpu_usage_pct, t.mem_usage_pct, 2 if ready_only else int(bool(exact)))
            t.names = [nd.name for nd in nodes]
            return t
        for i, nd in enumerate(nodes):
            t.alloc_cpu_milli[i] = 0 if ready_only else cores_to_milli(nd.available_cpu, exact)
            t.alloc_mem_bytes[i] = 0 if ready_only else gib_to_bytes(nd.available_memory, exact)
            t.pod_count[i] = nd.pod_count
            t.max_pods[i] = nd.max_pods
            t.ready[i] = 1 if is_ready(nd.conditions) else 0
            t.cpu_usage_pct[i] = nd.cpu_usage_percent
            t.mem_usage_pct[i] = nd.memory_usage_percent
            t.names[i] = nd.name
        return t

    @classmethod
    def from_allocatable(cls, rows: Sequence[Dict[str, Any]]) -> "NodeTable":
        """From K8s-style rows {name, cpu, memory, pods, pod_count, conditions}: the
        arithmetic of get_node_metrics (scheduler.py:139-151) without the REST calls."""
        recs = []
        for r in rows:
            mp = int(r["pods"])
            u = usage_percent(int(r["pod_count"]), mp)
            recs.append(NodeMetrics(r["name"], u, u, parse_cpu(r["cpu"]), parse_memory(r["memory"]),
                                    int(r["pod_count"]), mp,
                                    conditions=r.get("conditions",
                                                     [{"type": "Ready", "status": "True"}])))
        return cls.from_node_metrics(recs)


@dataclass
class PodTable:
    req_cpu_milli: np.ndarray
    req_mem_bytes: np.ndarray
    strategy_ids: Optional[np.ndarray] = None
    names: Optional[List[str]] = None

    def __len__(self) -> int:
        return len(self.req_cpu_milli)

    @classmethod
    def from_columns(cls, req_cpu_milli, req_mem_bytes, strategy_ids=None, names=None) -> "PodTable":
        return cls(np.ascontiguousarray(req_cpu_milli, dtype=np.int64),
                   np.ascontiguousarray(req_mem_bytes, dtype=np.int64),
                   None if strategy_ids is None else np.ascontiguousarray(strategy_ids, dtype=np.int64),
                   list(names) if names is not None else None)

    @classmethod
    def from_pod_specs(cls, pods: Sequence[Any], strategies: Optional[Sequence[str]] = None,
                       exact: bool = True) -> "PodTable":
        walker = _record_walker()
        if walker is not None:
            cpu, mem = np.empty(len(pods), np.int64), np.empty(len(pods), np.int64)
            walker.pod_specs_to_columns(pods, cpu, mem, exact)
        else:
            cpu = np.array([cores_to_milli(p.cpu_request, exact) for p in pods], dtype=np.int64)
            mem = np.array([gib_to_bytes(p.memory_request, exact) for p in pods], dtype=np.int64)
        sid = None if strategies is None else np.array([strategy_id(s) for s in strategies], np.int64)
        return cls(cpu, mem, sid, [p.name for p in pods])

    @classmethod
    def from_container_requests(cls, pods: Sequence[Sequence[Optional[Dict[str, Any]]]],
                                exact: bool = True) -> "PodTable":
        """From per-pod lists of container ``requests`` mappings (scheduler.py:731-753)."""
        specs = []
        for j, containers in enumerate(pods):
            c, m = pod_requests(containers)
            specs.append(PodSpec(f"pod-{j}", "default", c, m))
        return cls.from_pod_specs(specs, exact=exact)
