"""Batched caller for the fallback path.

The reference handles one watch event at a time: ``_watch_pods`` filters the event
(scheduler.py:673-675) and ``_schedule_pod`` (:690-729) converts the pod, re-lists the nodes,
decides, binds, and only then looks at the next event -- so every decision sees the binds before
it.  ``BatchScheduler.schedule_batch`` gives the same decisions, binds and statistics for a whole
list of pending pods with ONE sequential-commit placement on the GPU (``kp_session_greedy``) in
arrival order.  The only thing a batch cannot know in advance is a bind that fails: the reference
would then re-list a node table without that pod, so the batch is re-placed from the first failed
bind onwards with the node state rebuilt from the binds that did succeed.

Only the non-LLM decision is batched; the LLM call, cache and circuit breaker stay in the
reference (out of scope).  Ownership and error behaviour follow the reference: inputs are
borrowed, nothing raises into the caller's loop except what ``_convert_pod_to_spec`` itself would
raise for a malformed quantity string, and "no node" is the in-band ``selected_node == ""``.
"""
from __future__ import annotations

import logging
from typing import Any, Callable, Dict, List, Mapping, Optional, Sequence, Tuple

import numpy as np

from ._lib import FILTER_READY
from .tables import NodeTable, PodTable, SchedulingDecision, pod_requests, strategy_id

logger = logging.getLogger(__name__)

GreedyFn = Callable[[NodeTable, PodTable, int, int], np.ndarray]


def container_requests(pod: Any) -> List[Optional[Dict[str, Any]]]:
    """The ``requests`` mapping of every container, None where the reference skips it (:738)."""
    out = []
    for c in (pod.spec.containers or []):
        res = getattr(c, "resources", None)
        out.append(res.requests if res and res.requests else None)
    return out


class BatchScheduler:
    def __init__(self, scheduler_name: str, config: Optional[Mapping[str, Any]] = None, device: int = 0,
                 max_nodes: int = 65536, max_pods: int = 1 << 20, filter_mode: int = FILTER_READY,
                 decision_cls: Callable[..., Any] = SchedulingDecision,
                 greedy_fn: Optional[GreedyFn] = None,
                 pod_table_fn: Optional[Callable[[Sequence[Sequence[Any]]], PodTable]] = None):
        self.scheduler_name = scheduler_name
        self.config = config if config is not None else {}
        self.filter_mode = filter_mode
        self.decision_cls = decision_cls
        self.stats = {"total_scheduled": 0, "llm_decisions": 0, "fallback_decisions": 0,
                      "failed_bindings": 0}                                   # scheduler.py:632-637
        self._session = None
        if greedy_fn is None:
            from .placement import Session                   # raises if the CUDA library is missing
            self._session = Session(device=device, max_nodes=max_nodes, max_pods=max_pods)
            greedy_fn = self._gpu_greedy
        if pod_table_fn is None:
            from .ingest import ingest_pod_requests
            # multi-container fp64 sums such as 0.1 + 0.2 cores are rounded to the nearest milli-core /
            # byte for the capacity filter (which is an extension; the Ready filter ignores requests)
            pod_table_fn = lambda reqs: ingest_pod_requests(reqs, device=f"cuda:{device}", exact=False).table  # noqa: E731
        self._greedy = greedy_fn
        self._pod_table = pod_table_fn

    def _gpu_greedy(self, nodes: NodeTable, pods: PodTable, strategy: int, filter_mode: int) -> np.ndarray:
        return self._session.greedy(nodes, pods, strategy, filter_mode)[0]

    def close(self) -> None:
        if self._session is not None:
            self._session.close()

    def strategy(self) -> str:
        return self.config.get("fallback", {}).get("strategy", "resource_balanced")   # scheduler.py:526

    def needs_scheduling(self, pod: Any) -> bool:
        """The event filter of ``_watch_pods`` (scheduler.py:673-675)."""
        return (pod.status.phase == "Pending" and pod.spec.scheduler_name == self.scheduler_name
                and pod.spec.node_name is None)

    def schedule_batch(self, pods: Sequence[Any], node_metrics: Sequence[Any],
                       bind: Callable[[str, str, str], bool],
                       reason: str = "batched fallback") -> List[Tuple[Any, Any, bool]]:
        """Decide and bind ``pods`` (arrival order) against ``node_metrics`` (the table
        ``get_node_metrics()`` returned when the batch was drained).  Returns
        ``(pod, decision, bound)`` per pod; updates ``self.stats`` like ``_schedule_pod``."""
        pods = list(pods)
        if not pods:
            return []
        if not node_metrics:                                                   # scheduler.py:694-696
            logger.error("No available nodes")
            return []
        strategy = self.strategy()
        sid = strategy_id(strategy)
        table0 = NodeTable.from_node_metrics(node_metrics, exact=False)
        requests = [container_requests(p) for p in pods]
        req_table = self._requests_table(requests)
        n = len(pods)
        chosen = np.full(n, -1, dtype=np.int64)
        bound = np.zeros(n, dtype=bool)
        successes = np.zeros(len(table0), dtype=np.int64)
        start = 0
        while start < n:
            state = self._state_after(table0, successes)
            sub = PodTable(req_table.req_cpu_milli[start:], req_table.req_mem_bytes[start:])
            idx = np.asarray(self._greedy(state, sub, sid, self.filter_mode))
            failed_at = -1
            for k in range(n - start):
                j = start + k
                chosen[j] = idx[k]
                self.stats["fallback_decisions"] += 1                          # scheduler.py:708-710
                if idx[k] < 0:
                    self.stats["failed_bindings"] += 1                         # :727-729
                    continue
                meta = pods[j].metadata
                if bind(meta.name, meta.namespace, node_metrics[int(idx[k])].name):   # :716-721
                    self.stats["total_scheduled"] += 1
                    bound[j] = True
                    successes[int(idx[k])] += 1
                else:
                    self.stats["failed_bindings"] += 1                         # :725
                    failed_at = j
                    break
            if failed_at < 0:
                break
            start = failed_at + 1              # the rest saw a commit that never happened: place again
        out = []
        for j, pod in enumerate(pods):
            i = int(chosen[j])
            if i >= 0:                                                         # scheduler.py:551-557
                dec = self.decision_cls(selected_node=node_metrics[i].name, confidence=0.4,
                                        reasoning=f"Fallback ({strategy}): {reason}", fallback_needed=True)
            else:                                                              # :558-559
                dec = self.decision_cls("", 0.0, f"Fallback failed: {reason}", True)
            out.append((pod, dec, bool(bound[j])))
        return out

    def _requests_table(self, requests: Sequence[Sequence[Any]]) -> PodTable:
        if self.filter_mode == FILTER_READY:
            # the reference's own filter never looks at the pod, but _convert_pod_to_spec runs first
            # and raises on a malformed quantity (scheduler.py:692): keep that behaviour
            for r in requests:
                pod_requests(r)
            z = np.zeros(len(requests), dtype=np.int64)
            return PodTable(z, z.copy())
        return self._pod_table(requests)

    @staticmethod
    def _state_after(table0: NodeTable, successes: np.ndarray) -> NodeTable:
        """Node table the reference would re-list after ``successes[i]`` binds to node i: the count
        grows, and a node that received a bind reports the usage formula of scheduler.py:150-151;
        untouched nodes keep the usage they came with."""
        if not successes.any():
            return table0
        pc = table0.pod_count + successes
        mp = table0.max_pods
        with np.errstate(divide="ignore", invalid="ignore"):
            a3 = np.where(mp > 0, (pc / np.where(mp > 0, mp, 1)) * 50, 0.0)
        touched = successes > 0
        return NodeTable(table0.alloc_cpu_milli, table0.alloc_mem_bytes, pc, mp, table0.ready,
                         np.where(touched, a3, table0.cpu_usage_pct), np.where(touched, a3, table0.mem_usage_pct),
                         table0.names)


__all__ = ["BatchScheduler", "container_requests"]
