"""Drop-in for the reference's ``HuggingFaceClient._fallback_decision`` (scheduler.py:521-559).

The reference has no plugin API: the boundary is that one private method plus the config
key ``fallback.strategy`` (scheduler.py:526, config.yaml:34-36), read on every call.
``FallbackPlacer.fallback_decision(node_metrics, reason)`` has the same arguments, the same
return record and the same error behaviour (it never raises; "no result" is the in-band
value ``selected_node == ""``, confidence 0.0), and ``install(scheduler_module)`` assigns
it over the method without touching the reference file.

The decision itself is computed on the GPU through the C ABI (kp_session_snapshot with
one pod, KP_FILTER_READY -- the reference's own filter).  There is no CPU path here: if the
CUDA library cannot be loaded the constructor raises.

A ``FallbackPlacer`` owns one session and one set of column buffers that every call reuses: like the
reference's scheduler object it serves one scheduling loop -- calls from several threads need a placer each.
"""
from __future__ import annotations

import logging
from typing import Any, Callable, List, Mapping, Optional, Sequence

import numpy as np

from ._lib import FILTER_CAPACITY, FILTER_READY, KpNodeTable
from .placement import Session
from .tables import NODE_COLUMNS, NodeTable, PodTable, SchedulingDecision, _record_walker, strategy_id

logger = logging.getLogger(__name__)


class _ColumnBuffers:
    """Node-table columns kept across calls: the C record walker (csrc/kp_pyobj.c) fills rows [0, n) in place and the
    ready-made KpNodeTable points at them -- a decision allocates no arrays and rebuilds no ctypes structure."""
    _DTYPES = {"alloc_cpu_milli": np.int64, "alloc_mem_bytes": np.int64, "pod_count": np.int64, "max_pods": np.int64,
               "ready": np.uint8, "cpu_usage_pct": np.float64, "mem_usage_pct": np.float64}

    def __init__(self, capacity: int):
        self.capacity = capacity
        self.arrays = [np.zeros(capacity, self._DTYPES[k]) for k in NODE_COLUMNS]
        self.struct = KpNodeTable()
        for k, a in zip(NODE_COLUMNS, self.arrays):
            setattr(self.struct, k, a.ctypes.data)

    def fill(self, walker, node_metrics, mode: int) -> KpNodeTable:
        walker.node_metrics_to_columns(node_metrics, *self.arrays, mode | 4)      # + 4: buffers larger than the list
        self.struct.n_nodes = len(node_metrics)
        return self.struct


class FallbackPlacer:
    def __init__(self, config: Optional[Mapping[str, Any]] = None, device: int = 0,
                 max_nodes: int = 65536, max_pods: int = 65536,
                 decision_cls: Callable[..., Any] = SchedulingDecision):
        # `config` is the reference's CONFIG mapping (a live reference, so a strategy flipped
        # at run time is seen by the next call, like scheduler.py:526)
        self.config = config if config is not None else {}
        self.decision_cls = decision_cls
        self.session = Session(device=device, max_nodes=max_nodes, max_pods=max_pods)
        # one-decision calls go through kp_session_decide (one kernel, no copies, no stream synchronisation for
        # small tables); the node columns are written into buffers kept across calls
        self._cols = _ColumnBuffers(min(max_nodes, 1 << 16))

    def strategy(self) -> str:
        return self.config.get("fallback", {}).get("strategy", "resource_balanced")

    def _node_columns(self, node_metrics, ready_only: bool):
        """The node table of one decision: the kept column buffers when the C walker is there and the list fits."""
        walker = _record_walker()
        if walker is not None and len(node_metrics) <= self._cols.capacity:
            return self._cols.fill(walker, node_metrics, 2 if ready_only else 1)
        return NodeTable.from_node_metrics(node_metrics, ready_only=ready_only)

    # ---- same shape as the reference method (minus self) ----
    def fallback_decision(self, node_metrics: Sequence[Any], reason: str):
        if not node_metrics:                                        # scheduler.py:523-524
            return self.decision_cls("", 0.0, "No nodes available", True)
        strategy = self.strategy()
        try:
            # Ready-only filter: available_cpu / available_memory are dead inputs (the reference never reads
            # them, scheduler.py:531-549) and are not even looked at
            i, _ = self.session.decide(self._node_columns(node_metrics, ready_only=True), strategy_id(strategy), FILTER_READY)
            if i >= 0:                                              # scheduler.py:551-557
                return self.decision_cls(selected_node=node_metrics[i].name, confidence=0.4,
                                         reasoning=f"Fallback ({strategy}): {reason}",
                                         fallback_needed=True)
        except Exception as e:  # noqa: BLE001 -- never raise into the scheduling loop
            logger.error("GPU fallback placement failed: %s", e)
        return self.decision_cls("", 0.0, f"Fallback failed: {reason}", True)   # :558-559

    # ---- extension: the same decision with the capacity filter for a given pod ----
    def fallback_decision_for_pod(self, node_metrics: Sequence[Any], pod_spec: Any, reason: str):
        if not node_metrics:
            return self.decision_cls("", 0.0, "No nodes available", True)
        strategy = self.strategy()
        try:
            pods = PodTable.from_pod_specs([pod_spec])
            i, _ = self.session.decide(self._node_columns(node_metrics, ready_only=False), strategy_id(strategy), FILTER_CAPACITY,
                                       int(pods.req_cpu_milli[0]), int(pods.req_mem_bytes[0]))
            if i >= 0:
                return self.decision_cls(selected_node=node_metrics[i].name, confidence=0.4,
                                         reasoning=f"Fallback ({strategy}): {reason}",
                                         fallback_needed=True)
        except Exception as e:  # noqa: BLE001 -- same contract as fallback_decision
            logger.error("GPU fallback placement failed: %s", e)
        return self.decision_cls("", 0.0, f"Fallback failed: {reason}", True)

    # ---- batched forms over reference-style records ----
    def place_snapshot(self, node_metrics: Sequence[Any], pod_specs: Sequence[Any],
                       filter_mode: int = FILTER_CAPACITY) -> List[str]:
        table = NodeTable.from_node_metrics(node_metrics)
        pods = PodTable.from_pod_specs(pod_specs)
        idx, _ = self.session.snapshot(table, pods, strategy_id(self.strategy()), filter_mode)
        return [node_metrics[i].name if i >= 0 else "" for i in idx.tolist()]

    def place_greedy(self, node_metrics: Sequence[Any], pod_specs: Sequence[Any],
                     filter_mode: int = FILTER_CAPACITY) -> List[str]:
        table = NodeTable.from_node_metrics(node_metrics)
        pods = PodTable.from_pod_specs(pod_specs)
        idx, *_ = self.session.greedy(table, pods, strategy_id(self.strategy()), filter_mode)
        return [node_metrics[i].name if i >= 0 else "" for i in idx.tolist()]

    def install(self, scheduler_module) -> None:
        """Replace ``HuggingFaceClient._fallback_decision`` of an imported reference module."""
        self.config = scheduler_module.CONFIG
        self.decision_cls = scheduler_module.SchedulingDecision
        placer = self

        def _fallback_decision(_self, node_metrics, reason):
            return placer.fallback_decision(node_metrics, reason)

        scheduler_module.HuggingFaceClient._fallback_decision = _fallback_decision

    def close(self):
        self.session.close()
