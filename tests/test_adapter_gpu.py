"""The drop-in adapter: same signature, return record and error behaviour as the reference's
HuggingFaceClient._fallback_decision (scheduler.py:521-559), decisions computed on the GPU."""
import json
import types

import pytest

from k8s_llm_scheduler_b200.tables import NodeMetrics, PodSpec, SchedulingDecision
from tests.helpers import COND, from_bits

pytestmark = pytest.mark.gpu
STRATS = ["resource_balanced", "least_loaded", "round_robin", "no_such_strategy"]


@pytest.fixture(scope="module")
def placer():
    from k8s_llm_scheduler_b200.adapter import FallbackPlacer
    p = FallbackPlacer(config={"fallback": {"strategy": "resource_balanced"}}, max_nodes=4096, max_pods=4096)
    yield p
    p.close()


def records(case):
    # the "odd-alloc" cases carry available_cpu / available_memory values (inf, nan, 1e300, negative) that
    # the reference accepts because its decision never reads those fields (scheduler.py:531-549)
    return [NodeMetrics(f"node-{i}", from_bits(n["cu_bits"]), from_bits(n["mu_bits"]),
                        from_bits(n["avail_cpu_bits"]) if "avail_cpu_bits" in n else 8.0,
                        from_bits(n["avail_mem_bits"]) if "avail_mem_bits" in n else 15.616012573242188,
                        n["pc"], n["mp"], conditions=COND[n["cond"]])
            for i, n in enumerate(case["nodes"])]


def test_reference_decisions_through_adapter(placer, golden_dir):
    cases = json.loads((golden_dir / "decisions.json").read_text())
    for case in cases:
        nodes = records(case)
        for s in STRATS:
            placer.config["fallback"]["strategy"] = s          # read per call, like scheduler.py:526
            exp = case["expect"][s]
            dec = placer.fallback_decision(nodes, f"r-{case['tag']}")
            assert isinstance(dec, SchedulingDecision)
            want = "" if exp["idx"] < 0 else nodes[exp["idx"]].name
            assert dec.selected_node == want, (case["tag"], s)
            assert dec.confidence == exp["confidence"]
            assert dec.reasoning == exp["reasoning"]
            assert dec.fallback_needed is True


def test_adapter_never_raises(placer, golden_dir):
    """no exception may reach the scheduling loop; anything the GPU path cannot take comes
    back as the reference's in-band failure record (scheduler.py:558-559)."""
    class Broken:
        name = "broken"
        conditions = COND["ready"]
        cpu_usage_percent = memory_usage_percent = 1.0
        available_cpu = available_memory = 1.0
        max_pods = 110

        @property
        def pod_count(self):
            raise RuntimeError("record went away")

    placer.config["fallback"] = {"strategy": "least_loaded"}
    for bad in ([Broken()], [NodeMetrics("x", 1.0, 1.0, 1.0, 1.0, "seven", 110, conditions=COND["ready"])],
                [NodeMetrics("y", 1.0, 1.0, 1.0, 1.0, 2 ** 70, 110, conditions=COND["ready"])], [object()]):
        dec = placer.fallback_decision(bad, "why")
        assert (dec.selected_node, dec.confidence, dec.reasoning, dec.fallback_needed) == ("", 0.0, "Fallback failed: why", True)
    ok = placer.fallback_decision([NodeMetrics("z", 1.0, 1.0, float("nan"), float("inf"), 3, 110, conditions=COND["ready"])], "why")
    assert ok.selected_node == "z" and ok.confidence == 0.4
    # the per-pod extension reads the allocatable fields, so there the same record is a failure, not an exception
    from k8s_llm_scheduler_b200.tables import PodSpec
    dec = placer.fallback_decision_for_pod([NodeMetrics("z", 1.0, 1.0, float("nan"), 2.0, 3, 110, conditions=COND["ready"])],
                                           PodSpec("p", "default", 0.1, 0.125), "why")
    assert dec.selected_node == "" and dec.reasoning == "Fallback failed: why"


def test_default_strategy_and_capacity_extension(placer):
    placer.config.clear()                                       # default = resource_balanced
    nodes = [NodeMetrics("a", 45.0, 45.0, 4.0, 16.0, 99, 110, conditions=COND["ready"]),
             NodeMetrics("b", 5.0, 5.0, 1.0, 2.0, 11, 110, conditions=COND["ready"]),
             NodeMetrics("c", 5.0, 5.0, 1.0, 2.0, 11, 110, conditions=COND["false"])]
    dec = placer.fallback_decision(nodes, "x")
    assert dec.selected_node == "b" and dec.reasoning == "Fallback (resource_balanced): x"
    big = PodSpec("p", "default", 2.0, 1.0)                     # 2 cores: only "a" has room
    assert placer.fallback_decision_for_pod(nodes, big, "x").selected_node == "a"
    huge = PodSpec("p", "default", 64.0, 1.0)
    dec = placer.fallback_decision_for_pod(nodes, huge, "why")
    assert (dec.selected_node, dec.confidence, dec.reasoning) == ("", 0.0, "Fallback failed: why")
    assert placer.place_snapshot(nodes, [big, huge, PodSpec("q", "d", 0.5, 0.5)]) == ["a", "", "b"]
    assert placer.place_greedy(nodes, [PodSpec("q", "d", 0.5, 0.5)] * 3) == ["b", "b", "b"]


def test_install_over_reference_module(placer):
    class HuggingFaceClient:
        def _fallback_decision(self, node_metrics, reason):
            raise AssertionError("should have been replaced")

    mod = types.SimpleNamespace(CONFIG={"fallback": {"strategy": "least_loaded"}},
                                SchedulingDecision=SchedulingDecision,
                                HuggingFaceClient=HuggingFaceClient)
    placer.install(mod)
    nodes = [NodeMetrics("x", 10.0, 10.0, 8.0, 16.0, 5, 110, conditions=COND["ready"]),
             NodeMetrics("y", 9.0, 9.0, 8.0, 16.0, 5, 110, conditions=COND["ready"])]
    dec = mod.HuggingFaceClient()._fallback_decision(nodes, "Circuit breaker open")
    assert dec.selected_node == "y" and dec.reasoning == "Fallback (least_loaded): Circuit breaker open"
    mod.CONFIG["fallback"]["strategy"] = "round_robin"
    assert mod.HuggingFaceClient._fallback_decision(None, nodes, "z").selected_node == "x"
    assert mod.HuggingFaceClient._fallback_decision(None, [], "z").reasoning == "No nodes available"
