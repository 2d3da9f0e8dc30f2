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


def test_c_record_walker_equals_the_python_loops(monkeypatch):
    """csrc/kp_pyobj.c against the Python loops of tables.py on the same records: columns, readiness
    (every condition variant of the golden decisions), exactness errors, range errors."""
    import k8s_llm_scheduler_b200.tables as T
    from k8s_llm_scheduler_b200.build import build_pyobj
    from tests.helpers import COND
    build_pyobj()
    rng = np.random.Generator(np.random.PCG64(17))
    conds = list(COND.values()) + [[{"type": "Ready"}], [{"status": "True"}], [{"type": "Ready", "status": True}]]
    nodes = [T.NodeMetrics(f"n{i}", float(rng.random() * 50), 0 if i % 7 == 0 else float(rng.random() * 50),
                           int(rng.integers(1, 64000)) / 1000.0, int(rng.integers(1, 1 << 40)) / 1024 / 1024 / 1024,
                           int(rng.integers(0, 200)), int(rng.integers(0, 200)), conditions=conds[i % len(conds)])
             for i in range(500)]
    pods = [T.PodSpec(f"p{j}", "d", int(rng.integers(0, 64000)) / 1000.0, int(rng.integers(0, 1 << 36)) / 2 ** 30)
            for j in range(500)]

    def both(fn):
        monkeypatch.setattr(T, "_WALKER", [False])
        monkeypatch.delenv("KP_NO_PYOBJ", raising=False)
        assert T._record_walker() is not None
        fast = fn()
        monkeypatch.setattr(T, "_WALKER", [None])
        return fast, fn()

    a, b = both(lambda: T.NodeTable.from_node_metrics(nodes))
    for col in T.NODE_COLUMNS:
        assert np.array_equal(getattr(a, col), getattr(b, col)), col
        assert getattr(a, col).dtype == getattr(b, col).dtype
    assert a.names == b.names and a.ready.sum() not in (0, len(nodes))
    a, b = both(lambda: T.PodTable.from_pod_specs(pods))
    assert np.array_equal(a.req_cpu_milli, b.req_cpu_milli) and np.array_equal(a.req_mem_bytes, b.req_mem_bytes)
    for bad in (0.1 + 0.2, float("nan"), 1e300):
        for exact in (True, False):
            def one(bad=bad, exact=exact):
                try:
                    return T.PodTable.from_pod_specs([T.PodSpec("p", "d", bad, 1.0)], exact=exact).req_cpu_milli.tolist()
                except ValueError:
                    return "ValueError"
            fast, slow = both(one)
            assert fast == slow, (bad, exact)
    # ready_only (the adapter's form for the reference's own filter): the allocatable fields are not read
    odd = [T.NodeMetrics(f"o{i}", 3.0, 4.0, v, w, 5, 110, conditions=COND["ready"])
           for i, (v, w) in enumerate([(float("inf"), float("nan")), (1e300, -4.0), ("text", None)])]
    a, b = both(lambda: T.NodeTable.from_node_metrics(odd, ready_only=True))
    for col in T.NODE_COLUMNS:
        assert np.array_equal(getattr(a, col), getattr(b, col)), col
    assert not a.alloc_cpu_milli.any() and not a.alloc_mem_bytes.any() and a.pod_count.tolist() == [5, 5, 5]
    for exact in (True, False):
        fast, slow = both(lambda exact=exact: _raises(lambda: T.NodeTable.from_node_metrics(odd[:2], exact=exact)))
        assert fast == slow == "ValueError"


def _raises(fn):
    try:
        fn()
    except ValueError:
        return "ValueError"
    return "ok"


def test_record_walker_fills_column_buffers_kept_across_calls():
    """mode + 4 of the C walker (the adapter's one-decision path): buffers larger than the list, rows [0, n) written,
    the rest untouched; without the flag the sizes must match exactly; shorter buffers are refused either way."""
    import k8s_llm_scheduler_b200.tables as T
    from k8s_llm_scheduler_b200.build import build_pyobj
    from tests.helpers import COND
    build_pyobj()
    from k8s_llm_scheduler_b200 import _kp_pyobj as walker
    nodes = [T.NodeMetrics(f"n{i}", 10.0 + i, 20.0 + i, 8.0, 16.0, i, 110, conditions=COND["ready" if i % 2 else "false"])
             for i in range(5)]
    dt = {"alloc_cpu_milli": np.int64, "alloc_mem_bytes": np.int64, "pod_count": np.int64, "max_pods": np.int64, "ready": np.uint8,
          "cpu_usage_pct": np.float64, "mem_usage_pct": np.float64}
    cap = [np.full(16, 77, dt[k]) for k in T.NODE_COLUMNS]
    walker.node_metrics_to_columns(nodes, *cap, 1 | 4)
    want = T.NodeTable.from_node_metrics(nodes)
    for k, buf in zip(T.NODE_COLUMNS, cap):
        assert np.array_equal(buf[:5], getattr(want, k)), k
        assert (buf[5:] == 77).all(), k
    walker.node_metrics_to_columns(nodes, *cap, 2 | 4)                      # ready-only: allocatable columns zeroed, not read
    assert not cap[0][:5].any() and not cap[1][:5].any() and cap[2][:5].tolist() == [0, 1, 2, 3, 4]
    with pytest.raises(ValueError):
        walker.node_metrics_to_columns(nodes, *cap, 1)                      # exact-size mode refuses larger buffers
    short = [np.zeros(3, dt[k]) for k in T.NODE_COLUMNS]
    with pytest.raises(ValueError):
        walker.node_metrics_to_columns(nodes, *short, 1 | 4)
