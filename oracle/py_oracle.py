"""CPU oracle for the fallback placement path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference
legs may import this module.  The product package (``k8s_llm_scheduler_b200``) never
does: it fails loudly when its CUDA library is missing instead of falling back here.

What this is: a restatement, in plain CPython, of the arithmetic the reference performs
on its non-LLM placement path.  Every function cites the reference lines it follows
(``scheduler.py`` = /root/reference/scheduler.py).  It deliberately keeps the reference's
data shapes (a list of per-node records with a ``conditions`` list of dicts, a float
score, a strict ``>`` against a ``-1`` sentinel) so that (i) results are bit-identical
and (ii) timing it is representative of timing the reference's own CPython loop.

Parity status: PINNED.  tests/test_oracle_golden.py checks this module against the
golden vectors in tests/golden/*.json, which tests/golden/make_golden.py produced by
running the unmodified reference in the build container (score bits captured from the
reference function's own locals).

Two filters are provided:

  FILTER_READY (0)     what the reference does: Ready nodes only (scheduler.py:531-535).
  FILTER_CAPACITY (1)  the north-star extension: additionally
                       pod_count < max_pods, cpu_req <= free_cpu and mem_req <= free_mem,
                       with "free" defined by the only capacity formula the reference
                       contains (scheduler.py:232-233).  Where every pod fits every Ready
                       node with room, both filters give the same answer (tested).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

FILTER_READY = 0
FILTER_CAPACITY = 1

STRATEGY_IDS = {"resource_balanced": 0, "least_loaded": 1}   # anything else -> 2 (scheduler.py:544)
STRATEGY_NAMES = {0: "resource_balanced", 1: "least_loaded", 2: "round_robin"}

READY_CONDITIONS = [{"type": "Ready", "status": "True", "reason": "KubeletReady"}]
NOT_READY_CONDITIONS = [{"type": "Ready", "status": "False", "reason": "KubeletNotReady"}]


# ------------------------------------------------------------------ records (scheduler.py:72-104)
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


# ------------------------------------------------------------------ quantity parsing
def parse_cpu(text: str) -> float:
    """K8s CPU quantity -> cores.  Follows scheduler.py:172-176: only the 'm' suffix."""
    if text.endswith("m"):
        return float(text[:-1]) / 1000
    return float(text)


def parse_memory(text: str) -> float:
    """K8s memory quantity -> GiB.  Follows scheduler.py:178-187: Ki / Mi / Gi, else bytes."""
    if text.endswith("Ki"):
        return float(text[:-2]) / 1024 / 1024
    if text.endswith("Mi"):
        return float(text[:-2]) / 1024
    if text.endswith("Gi"):
        return float(text[:-2])
    return float(text) / 1024 / 1024 / 1024


def pod_requests(containers: Iterable[Optional[Dict[str, Any]]]) -> Tuple[float, float]:
    """Sum of container requests as the reference sums them (scheduler.py:737-753).

    ``containers`` holds one ``requests`` mapping (or None / empty) per container.  CPU:
    'm' strings are /1000, anything else ``float(x or 0)``.  Memory: ONLY Ki/Mi/Gi
    strings count; plain-byte strings and non-strings add nothing (differs from
    parse_memory -- a quirk of the reference that is kept).
    """
    cpu_total, mem_total = 0.0, 0.0
    for req in containers:
        if not req:
            continue
        cpu_s = req.get("cpu", "0")
        mem_s = req.get("memory", "0")
        if isinstance(cpu_s, str) and cpu_s.endswith("m"):
            cpu_total += float(cpu_s[:-1]) / 1000
        else:
            cpu_total += float(cpu_s or 0)
        if isinstance(mem_s, str):
            if mem_s.endswith("Ki"):
                mem_total += float(mem_s[:-2]) / 1024 / 1024
            elif mem_s.endswith("Mi"):
                mem_total += float(mem_s[:-2]) / 1024
            elif mem_s.endswith("Gi"):
                mem_total += float(mem_s[:-2])
    return cpu_total, mem_total


def usage_percent(pod_count: int, max_pods: int):
    """The reference's stand-in for metrics-server (scheduler.py:150-151)."""
    return (pod_count / max_pods) * 50 if max_pods > 0 else 0


def node_row(name: str, cpu: str, memory: str, pods: str, pod_count: int,
             conditions: Optional[List[Dict[str, str]]] = None) -> NodeMetrics:
    """One node-table row from allocatable strings + running-pod count (scheduler.py:139-165)."""
    max_pods = int(pods)
    u = usage_percent(pod_count, max_pods)
    return NodeMetrics(name=name, cpu_usage_percent=u, memory_usage_percent=u,
                       available_cpu=parse_cpu(cpu), available_memory=parse_memory(memory),
                       pod_count=pod_count, max_pods=max_pods,
                       conditions=READY_CONDITIONS if conditions is None else conditions)


# ------------------------------------------------------------------ the hot path
def node_is_ready(node: NodeMetrics) -> bool:
    """scheduler.py:531-535."""
    return any(c.get("type") == "Ready" and c.get("status") == "True" for c in node.conditions)


def node_score(node: NodeMetrics, strategy: str):
    """Per-node score, scheduler.py:537-545.  Operation order is part of the contract."""
    if strategy == "resource_balanced":
        cpu_score = (100 - node.cpu_usage_percent) / 100.0
        memory_score = (100 - node.memory_usage_percent) / 100.0
        pod_score = (node.max_pods - node.pod_count) / node.max_pods if node.max_pods > 0 else 0
        return (cpu_score * 0.35) + (memory_score * 0.35) + (pod_score * 0.30)
    if strategy == "least_loaded":
        return (100 - node.cpu_usage_percent) + (100 - node.memory_usage_percent)
    return node.pod_count          # round_robin and every unknown string (scheduler.py:544-545)


def free_capacity(node: NodeMetrics) -> Tuple[float, float]:
    """Free cores / GiB as the reference defines them (scheduler.py:232-233)."""
    free_cpu = node.available_cpu * (100 - node.cpu_usage_percent) / 100.0
    free_mem = node.available_memory * (100 - node.memory_usage_percent) / 100.0
    return free_cpu, free_mem


def pod_fits(pod: PodSpec, node: NodeMetrics) -> bool:
    """North-star capacity filter; no counterpart on the reference's
    fallback path -- defined here, and by construction equal to "always true" when the
    pod fits everywhere."""
    if not node.pod_count < node.max_pods:
        return False
    free_cpu, free_mem = free_capacity(node)
    return pod.cpu_request <= free_cpu and pod.memory_request <= free_mem


def pick_node(nodes: Sequence[NodeMetrics], strategy: str, pod: Optional[PodSpec] = None,
              filter_mode: int = FILTER_READY):
    """Filter -> score -> first-best argmax (scheduler.py:528-549).

    Returns (index, score): index -1 when nothing wins, score = the reference's
    ``best_score`` local at return (the -1 sentinel when nothing wins).
    """
    best_idx = -1
    best_score = -1
    for i, node in enumerate(nodes):
        if not node_is_ready(node):
            continue
        if filter_mode == FILTER_CAPACITY and not pod_fits(pod, node):
            continue
        score = node_score(node, strategy)
        if score > best_score:           # strict: first node in list order keeps ties
            best_score = score
            best_idx = i
    return best_idx, best_score


def fallback_decision(nodes: Sequence[NodeMetrics], reason: str, strategy: str,
                      pod: Optional[PodSpec] = None,
                      filter_mode: int = FILTER_READY) -> SchedulingDecision:
    """The reference's result construction (scheduler.py:523-524, 551-559)."""
    if not nodes:
        return SchedulingDecision("", 0.0, "No nodes available", True)
    idx, _ = pick_node(nodes, strategy, pod, filter_mode)
    if idx >= 0:
        return SchedulingDecision(nodes[idx].name, 0.4, f"Fallback ({strategy}): {reason}", True)
    return SchedulingDecision("", 0.0, f"Fallback failed: {reason}", True)


def commit(node: NodeMetrics) -> None:
    """State transition of one bind as the next get_node_metrics() sees it
    (scheduler.py:598 -> :144-151): one more pod, both usages re-derived from the count."""
    node.pod_count += 1
    u = usage_percent(node.pod_count, node.max_pods)
    node.cpu_usage_percent = u
    node.memory_usage_percent = u


def snapshot_place(nodes: Sequence[NodeMetrics], pods: Sequence[PodSpec],
                   strategies, filter_mode: int = FILTER_READY):
    """Every pod against the same, unchanging node table.  ``strategies``: one name or
    one name per pod.  Returns (indices, scores)."""
    one = isinstance(strategies, str)
    idxs, scores = [], []
    for j, pod in enumerate(pods):
        i, s = pick_node(nodes, strategies if one else strategies[j], pod, filter_mode)
        idxs.append(i)
        scores.append(float(s))
    return idxs, scores


def greedy_place(nodes: Sequence[NodeMetrics], pods: Sequence[PodSpec],
                 strategies, filter_mode: int = FILTER_READY):
    """Pods in arrival order, each bind visible to the next decision
    (scheduler.py:690-729).  MUTATES ``nodes``.  Returns (indices, scores)."""
    one = isinstance(strategies, str)
    idxs, scores = [], []
    for j, pod in enumerate(pods):
        i, s = pick_node(nodes, strategies if one else strategies[j], pod, filter_mode)
        idxs.append(i)
        scores.append(float(s))
        if i >= 0:
            commit(nodes[i])
    return idxs, scores


# ------------------------------------------------------------------ tensor form <-> records
def nodes_from_arrays(alloc_cpu_milli, alloc_mem_bytes, pod_count, max_pods, ready,
                      cpu_usage_pct, mem_usage_pct) -> List[NodeMetrics]:
    """int64 milli-CPU / bytes / counts + f64 usage -> records.  cores = milli/1000.0 and
    GiB = bytes/1024/1024/1024 are what parse_cpu / parse_memory give for the same
    quantity written as '<milli>m' / '<bytes>' (SURVEY.md C2)."""
    out = []
    for i in range(len(pod_count)):
        out.append(NodeMetrics(
            name=f"node-{i}", cpu_usage_percent=float(cpu_usage_pct[i]),
            memory_usage_percent=float(mem_usage_pct[i]),
            available_cpu=int(alloc_cpu_milli[i]) / 1000.0,
            available_memory=int(alloc_mem_bytes[i]) / 1024 / 1024 / 1024,
            pod_count=int(pod_count[i]), max_pods=int(max_pods[i]),
            conditions=READY_CONDITIONS if ready[i] else NOT_READY_CONDITIONS))
    return out


def pods_from_arrays(req_cpu_milli, req_mem_bytes) -> List[PodSpec]:
    return [PodSpec(name=f"pod-{j}", namespace="default",
                    cpu_request=int(req_cpu_milli[j]) / 1000.0,
                    memory_request=int(req_mem_bytes[j]) / 1024 / 1024 / 1024)
            for j in range(len(req_cpu_milli))]


def strategy_name(sid: int) -> str:
    return STRATEGY_NAMES.get(int(sid), "round_robin")
