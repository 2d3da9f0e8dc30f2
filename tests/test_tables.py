"""Host-side table builders (reference A1-A5) against the reference-generated vectors."""
import json

import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth, tables
from tests.helpers import COND, bits


def load(golden_dir, name):
    return json.loads((golden_dir / name).read_text())


def test_parsers_match_reference(golden_dir):
    g = load(golden_dir, "parsers.json")
    for r in g["parse_cpu"]:
        assert bits(tables.parse_cpu(r["in"])) == r["bits"]
    for r in g["parse_memory"]:
        assert bits(tables.parse_memory(r["in"])) == r["bits"]
    for r in g["parse_memory_errors"]:
        with pytest.raises(ValueError):
            tables.parse_memory(r["in"])
    for r in g["convert_pod_to_spec"]:
        reqs = [{k: v for k, v in (("cpu", c), ("memory", m)) if v is not None}
                for c, m in r["containers"]]
        cpu, mem = tables.pod_requests(reqs)
        assert bits(cpu) == r["cpu_bits"] and bits(mem) == r["mem_bits"]


def test_node_table_from_allocatable_matches_reference_rows(golden_dir):
    for case in load(golden_dir, "greedy.json"):
        rows = [dict(name=ns["name"], cpu=ns["cpu"], memory=ns["memory"], pods=ns["pods"],
                     pod_count=ns["n_running"], conditions=COND[ns.get("cond", "ready")])
                for ns in case["nodes"]]
        t = tables.NodeTable.from_allocatable(rows)
        for i, p in enumerate(case["parsed"]):
            assert bits(t.alloc_cpu_milli[i] / 1000.0) == p["available_cpu_bits"]
            assert bits(t.alloc_mem_bytes[i] / 1024 / 1024 / 1024) == p["available_memory_bits"]
            assert bits(t.cpu_usage_pct[i]) == p["cu_bits"]
            assert (t.max_pods[i], t.pod_count[i]) == (p["max_pods"], p["pod_count"])
            assert t.ready[i] == int(ns_ready(case["nodes"][i]))


def ns_ready(ns):
    return tables.is_ready(COND[ns.get("cond", "ready")])


def test_wire_format_exactness():
    assert tables.cores_to_milli(0.25) == 250 and tables.cores_to_milli(7.91) == 7910
    assert tables.gib_to_bytes(15.616012573242188) == 16374576 * 1024
    with pytest.raises(ValueError):
        tables.cores_to_milli(0.1 + 0.2)           # 0.30000000000000004 is not k/1000
    assert tables.cores_to_milli(0.1 + 0.2, exact=False) == 300
    assert tables.strategy_id("resource_balanced") == 0 and tables.strategy_id("least_loaded") == 1
    assert tables.strategy_id("round_robin") == 2 and tables.strategy_id("whatever") == 2


def test_pod_table_quirks():
    # plain-byte memory strings are ignored for pods (scheduler.py:747-753), multi-container sums
    t = tables.PodTable.from_container_requests([[{"cpu": "100m", "memory": "1073741824"}],
                                                 [{"cpu": "250m", "memory": "256Mi"}],
                                                 [{"cpu": "1", "memory": "1Gi"}, {"cpu": "500m", "memory": "512Mi"}]])
    assert t.req_cpu_milli.tolist() == [100, 250, 1500]
    assert t.req_mem_bytes.tolist() == [0, 256 << 20, (1 << 30) + (512 << 20)]


def test_synth_usage_matches_python_and_is_deterministic():
    n1, p1 = synth.config5(3000, 5000)
    n2, p2 = synth.config5(3000, 5000)
    for k in tables.NODE_COLUMNS:
        assert np.array_equal(getattr(n1, k), getattr(n2, k))
    assert np.array_equal(p1.req_cpu_milli, p2.req_cpu_milli)
    for i in range(0, 3000, 7):
        u = float(tables.usage_percent(int(n1.pod_count[i]), int(n1.max_pods[i])))
        assert bits(n1.cpu_usage_pct[i]) == bits(u)
    assert 0.015 < 1 - n1.ready.mean() < 0.03 and (n1.max_pods == 0).mean() < 0.02
    nodes, pods = synth.config1()
    assert pods.req_cpu_milli.tolist() == [250, 500, 100]
    assert pods.req_mem_bytes.tolist() == [256 << 20, 512 << 20, 128 << 20]
    assert bits(nodes.cpu_usage_pct[0]) == bits(3.6363636363636362)


def test_config5_infeasible_fraction():
    from oracle import c_oracle
    nodes, pods = synth.config5(2000, 4000)
    idx, _ = c_oracle.snapshot(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes,
                               strategy_ids=pods.strategy_ids, filter_mode=1, threads=4)
    frac = (idx < 0).mean()
    assert 0.27 < frac < 0.33, frac
