"""Prompt-side reuse of the hoisted per-node values.

``PromptEngine.construct_scheduling_prompt`` (scheduler.py:216-252) recomputes, for every pod, the
same per-node free capacity the placement kernels hoist (``available * (100 - usage%) / 100.0``,
:232-233) and prints it for EVERY node, so the prompt grows with the cluster.  This module renders
that node block from ``kp_node_scores`` (one kernel for the table, text byte-identical to the
reference's for the same nodes) and can cut the list down to the ``top_k`` nodes the pod fits, in
the fallback strategy's order -- the list the LLM is asked to choose from stays in cluster order.

Only the computed block is produced here; the surrounding prompt text stays in the reference.
"""
from __future__ import annotations

from typing import Any, List, Optional, Sequence

import numpy as np

from ._lib import FILTER_CAPACITY
from .tables import NodeTable, strategy_id


def _node_values(node_metrics: Sequence[Any], strategy: str, device: int):
    from . import placement                       # raises if the CUDA library is missing
    table = NodeTable.from_node_metrics(node_metrics)
    dn = placement.nodes_to_device(table, f"cuda:{device}")
    score, fc, fm, el = placement.node_scores(dn, strategy_id(strategy), FILTER_CAPACITY)
    return score.cpu().numpy(), fc.cpu().numpy(), fm.cpu().numpy(), el.cpu().numpy().astype(bool)


def cluster_info(node_metrics: Sequence[Any], device: int = 0, keep: Optional[Sequence[int]] = None) -> str:
    """The "AVAILABLE NODES" block and the "VALID NODE NAMES" line of the reference prompt
    (scheduler.py:226-243) for ``node_metrics`` (or for the rows listed in ``keep``)."""
    _, fc, fm, _ = _node_values(node_metrics, "resource_balanced", device) if node_metrics else (None, [], [], None)
    rows = range(len(node_metrics)) if keep is None else keep
    text = "AVAILABLE NODES:\n"
    names: List[str] = []
    for i in rows:
        node = node_metrics[i]
        names.append(node.name)
        text += f"""
{node.name}:
  - CPU Usage: {node.cpu_usage_percent:.1f}% (available: {fc[i]:.2f} cores)
  - Memory Usage: {node.memory_usage_percent:.1f}% (available: {fm[i]:.2f} GB)
  - Pods: {node.pod_count}/{node.max_pods}
  - Status: Ready
"""
    text += f"\nVALID NODE NAMES: {', '.join(names)}\n"
    return text


def prefilter(node_metrics: Sequence[Any], pod_spec: Any, strategy: str = "resource_balanced", top_k: int = 16,
              device: int = 0) -> List[int]:
    """Rows of the at most ``top_k`` nodes the pod fits (Ready, room for a pod, enough free cpu and
    memory), best first under the fallback strategy's score with ties to the earlier node, returned
    in cluster order so that a prompt built from them keeps the reference's node order."""
    if not node_metrics:
        return []
    score, fc, fm, el = _node_values(node_metrics, strategy, device)
    fits = el & (pod_spec.cpu_request <= fc) & (pod_spec.memory_request <= fm)
    rows = np.nonzero(fits)[0]
    order = rows[np.lexsort((rows, -score[rows]))][:top_k]      # score desc, index asc
    return sorted(int(i) for i in order)


__all__ = ["cluster_info", "prefilter"]
