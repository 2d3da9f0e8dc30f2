"""Prompt-side reuse (k8s_llm_scheduler_b200/prompt.py): the node block rendered from kp_node_scores is
byte-identical to what the reference's PromptEngine prints (tests/golden/prompts.json, produced by the
unmodified reference, scheduler.py:226-243), and the pre-filter keeps exactly the nodes the oracle says
the pod fits, best first."""
import json
from pathlib import Path

import numpy as np
import pytest

from k8s_llm_scheduler_b200.tables import NodeMetrics, PodSpec
from oracle import py_oracle
from tests.helpers import from_bits

pytestmark = pytest.mark.gpu
GOLD = json.loads((Path(__file__).parent / "golden" / "prompts.json").read_text())
READY = [{"type": "Ready", "status": "True"}]


def records(nodes):
    return [NodeMetrics(n["name"], from_bits(n["cu_bits"]), from_bits(n["cu_bits"]), n["cpu"], n["mem"],
                        n["pc"], n["mp"], conditions=READY) for n in nodes]


def test_node_block_is_byte_identical_to_the_reference():
    from k8s_llm_scheduler_b200.prompt import cluster_info
    assert len(GOLD) >= 4
    for case in GOLD:
        assert cluster_info(records(case["nodes"])) == case["cluster_info"]
    keep = [0]
    part = cluster_info(records(GOLD[-1]["nodes"]), keep=keep)
    assert part.count("  - Status: Ready") == 1 and GOLD[-1]["nodes"][0]["name"] in part


def test_prefilter_keeps_the_best_fitting_nodes():
    from k8s_llm_scheduler_b200.prompt import prefilter
    rng = np.random.Generator(np.random.PCG64(3))
    nodes = []
    for i in range(400):
        cpu, mem, mp = [("4", "16Gi", "30"), ("16", "64Gi", "58"), ("32", "128Gi", "110")][int(rng.integers(0, 3))]
        nodes.append(py_oracle.node_row(f"n{i}", cpu, mem, mp, int(rng.integers(0, int(mp) + 1)),
                                        READY if rng.random() > 0.1 else [{"type": "Ready", "status": "False"}]))
    for strategy in ("resource_balanced", "least_loaded", "round_robin"):
        for req in ((0.25, 0.25), (3.0, 20.0), (14.0, 100.0), (500.0, 1.0)):
            pod = PodSpec("p", "default", *req)
            opod = py_oracle.PodSpec("p", "default", *req)
            fit = [i for i, nd in enumerate(nodes) if py_oracle.node_is_ready(nd) and py_oracle.pod_fits(opod, nd)]
            best = sorted(fit, key=lambda i: (-py_oracle.node_score(nodes[i], strategy), i))[:16]
            assert prefilter(nodes, pod, strategy, top_k=16) == sorted(best), (strategy, req)
    assert prefilter([], PodSpec("p", "default", 1.0, 1.0)) == []
