"""Pins the CPU oracle (oracle/py_oracle.py, oracle/c_oracle.c) to the golden vectors
that tests/golden/make_golden.py produced by running the unmodified reference
(scheduler.py:521-559, :121-187, :731-764).  Bit-exact everywhere."""
import json
import struct

import numpy as np
import pytest

from oracle import c_oracle, py_oracle as po


def bits(x) -> str:
    return struct.pack(">d", float(x)).hex()


def from_bits(h: str) -> float:
    return struct.unpack(">d", bytes.fromhex(h))[0]


CONDS = {
    "ready": po.READY_CONDITIONS,
    "false": [{"type": "Ready", "status": "False", "reason": "x"}],
    "unknown": [{"type": "Ready", "status": "Unknown", "reason": "x"}],
    "none": [],
    "other_only": [{"type": "MemoryPressure", "status": "True", "reason": "x"}],
    "ready_second": [{"type": "DiskPressure", "status": "False", "reason": "x"},
                     {"type": "Ready", "status": "True", "reason": "x"}],
}
STRATS = ["resource_balanced", "least_loaded", "round_robin", "no_such_strategy"]


def load(golden_dir, name):
    return json.loads((golden_dir / name).read_text())


def test_kat_scores_a3(golden_dir):
    rows = load(golden_dir, "kat_scores.json")["a3_usage"]
    assert len(rows) > 200
    for r in rows:
        pc, mp = r["pc"], r["mp"]
        u = po.usage_percent(pc, mp)
        assert bits(u) == r["usage_bits"]
        assert bits(c_oracle.usage(pc, mp)) == r["usage_bits"]
        node = po.NodeMetrics("n", u, u, 8.0, 15.6, pc, mp, conditions=po.READY_CONDITIONS)
        for sid, s in enumerate(STRATS[:3]):
            idx, score = po.pick_node([node], s)
            assert (idx == 0) == r[s]["selected"], (pc, mp, s)
            assert bits(score) == r[s]["score_bits"], (pc, mp, s)
            assert isinstance(score, int) == r[s]["score_is_int"]
            if r[s]["selected"]:
                assert bits(c_oracle.score(sid, u, u, pc, mp)) == r[s]["score_bits"]


def test_kat_scores_survey_table():
    # SURVEY.md §8(c) known-answer table, typed in independently of the JSON
    table = [(2, 110, "3fef9f2f79f2f79f", "4068c5d1745d1746"),
             (3, 110, "3fef6ec736ec736e", "4068a8ba2e8ba2e9"),
             (8, 110, "3fee7cbde7cbde7c", "40681745d1745d17"),
             (55, 110, "3fe5999999999999", "4062c00000000000"),
             (1, 3, "3fe9111111111111", "4064d55555555556"),
             (0, 0, "3fe6666666666666", "4069000000000000"),
             (109, 110, "3fd6c736ec736ec7", "40593a2e8ba2e8ba"),
             (110, 110, "3fd6666666666666", "4059000000000000")]
    for pc, mp, rb, ll in table:
        u = float(po.usage_percent(pc, mp))
        assert bits(c_oracle.score(0, u, u, pc, mp)) == rb
        assert bits(c_oracle.score(1, u, u, pc, mp)) == ll
        node = po.NodeMetrics("n", u, u, 1.0, 1.0, pc, mp)
        assert bits(po.node_score(node, "resource_balanced")) == rb
        assert bits(po.node_score(node, "least_loaded")) == ll


def test_kat_scores_custom_usage(golden_dir):
    for r in load(golden_dir, "kat_scores.json")["custom_usage"]:
        node = po.NodeMetrics("n", r["cu"], r["mu"], 8.0, 15.6, r["pc"], r["mp"],
                              conditions=po.READY_CONDITIONS)
        for sid, s in enumerate(STRATS[:3]):
            idx, score = po.pick_node([node], s)
            assert (idx == 0) == r[s]["selected"]
            assert bits(score) == r[s]["score_bits"]
            got_idx, got_score = c_oracle.snapshot(
                dict(alloc_cpu_milli=[8000], alloc_mem_bytes=[1 << 34], pod_count=[r["pc"]],
                     max_pods=[r["mp"]], ready=[1], cpu_usage_pct=[r["cu"]],
                     mem_usage_pct=[r["mu"]]), [0], [0], strategy=sid, filter_mode=0)
            assert (got_idx[0] == 0) == r[s]["selected"]
            assert bits(got_score[0]) == r[s]["score_bits"]


def _case_nodes(case):
    nodes = [po.NodeMetrics(f"node-{i}", from_bits(n["cu_bits"]), from_bits(n["mu_bits"]),
                            8.0, 15.616012573242188, n["pc"], n["mp"], conditions=CONDS[n["cond"]])
             for i, n in enumerate(case["nodes"])]
    arrays = dict(alloc_cpu_milli=[8000] * len(nodes), alloc_mem_bytes=[16374576 * 1024] * len(nodes),
                  pod_count=[n.pod_count for n in nodes], max_pods=[n.max_pods for n in nodes],
                  ready=[int(po.node_is_ready(n)) for n in nodes],
                  cpu_usage_pct=[n.cpu_usage_percent for n in nodes],
                  mem_usage_pct=[n.memory_usage_percent for n in nodes])
    return nodes, arrays


def test_decisions(golden_dir):
    cases = load(golden_dir, "decisions.json")
    assert len(cases) >= 60
    for case in cases:
        nodes, arrays = _case_nodes(case)
        for s in STRATS:
            exp = case["expect"][s]
            dec = po.fallback_decision(nodes, f"r-{case['tag']}", s)
            assert dec.reasoning == exp["reasoning"]
            assert dec.confidence == exp["confidence"]
            assert dec.fallback_needed is exp["fallback_needed"]
            if not nodes:
                assert dec.selected_node == ""
                continue
            idx, score = po.pick_node(nodes, s)
            assert idx == exp["idx"], (case["tag"], s)
            assert bits(score) == exp["score_bits"]
            assert dec.selected_node == ("" if idx < 0 else nodes[idx].name)
            sid = po.STRATEGY_IDS.get(s, 2)
            ci, cs = c_oracle.snapshot(arrays, [100], [1 << 20], strategy=sid, filter_mode=0)
            assert ci[0] == exp["idx"], (case["tag"], s)
            assert bits(cs[0]) == exp["score_bits"]


def test_greedy_sequences(golden_dir):
    cases = load(golden_dir, "greedy.json")
    assert len(cases) >= 40
    for case in cases:
        nodes = [po.node_row(ns["name"], ns["cpu"], ns["memory"], ns["pods"], ns["n_running"],
                             CONDS[ns.get("cond", "ready")]) for ns in case["nodes"]]
        for n, p in zip(nodes, case["parsed"]):
            assert bits(n.available_cpu) == p["available_cpu_bits"]
            assert bits(n.available_memory) == p["available_memory_bits"]
            assert bits(n.cpu_usage_percent) == p["cu_bits"]
            assert (n.max_pods, n.pod_count) == (p["max_pods"], p["pod_count"])
        arrays = dict(alloc_cpu_milli=[round(n.available_cpu * 1000) for n in nodes],
                      alloc_mem_bytes=[round(n.available_memory * 2 ** 30) for n in nodes],
                      pod_count=[n.pod_count for n in nodes], max_pods=[n.max_pods for n in nodes],
                      ready=[int(po.node_is_ready(n)) for n in nodes],
                      cpu_usage_pct=[n.cpu_usage_percent for n in nodes],
                      mem_usage_pct=[n.memory_usage_percent for n in nodes])
        pods = [po.PodSpec(f"p{j}", "default", 0.1, 0.1) for j in range(case["n_pods"])]
        idxs, scores = po.greedy_place(nodes, pods, case["strategy"])
        assert idxs == case["idx_seq"], case["tag"]
        assert [bits(s) for s in scores] == case["score_bits_seq"], case["tag"]
        assert [n.pod_count for n in nodes] == case["final_pod_count"]
        sid = po.STRATEGY_IDS.get(case["strategy"], 2)
        ci, cs, pc, cu, mu = c_oracle.greedy(arrays, [100] * case["n_pods"],
                                             [1 << 20] * case["n_pods"], strategy=sid,
                                             filter_mode=0)
        assert ci.tolist() == case["idx_seq"], case["tag"]
        assert [bits(s) for s in cs] == case["score_bits_seq"]
        assert pc.tolist() == case["final_pod_count"]
        assert [bits(x) for x in cu] == [bits(n.cpu_usage_percent) for n in nodes]


def test_parsers(golden_dir):
    g = load(golden_dir, "parsers.json")
    for r in g["parse_cpu"]:
        assert bits(po.parse_cpu(r["in"])) == r["bits"]
    for r in g["parse_memory"]:
        assert bits(po.parse_memory(r["in"])) == r["bits"]
    for r in g["parse_memory_errors"]:
        assert r["error"] == "ValueError"
        with pytest.raises(ValueError):
            po.parse_memory(r["in"])
    for r in g["convert_pod_to_spec"]:
        reqs = []
        for cpu, mem in r["containers"]:
            d = {}
            if cpu is not None:
                d["cpu"] = cpu
            if mem is not None:
                d["memory"] = mem
            reqs.append(d)
        cpu, mem = po.pod_requests(reqs)
        assert bits(cpu) == r["cpu_bits"] and bits(mem) == r["mem_bits"]


def test_config1_trace(golden_dir):
    """BASELINE config 1: 3-node cluster x ai-test-pods.yaml; binds [m02, m03, m02]."""
    g = load(golden_dir, "config1_trace.json")
    assert g["binds"] == ["minikube-m02", "minikube-m03", "minikube-m02"]
    nodes = [po.node_row("minikube", "8", "16374576Ki", "110", 8),
             po.node_row("minikube-m02", "8", "16374576Ki", "110", 2),
             po.node_row("minikube-m03", "8", "16374576Ki", "110", 2)]
    pods = []
    for p in g["pods"]:
        cpu, mem = po.pod_requests([p["requests"]])
        assert bits(cpu) == p["cpu_bits"] and bits(mem) == p["mem_bits"]
        pods.append(po.PodSpec(p["name"], "default", cpu, mem))
    for mode in (po.FILTER_READY, po.FILTER_CAPACITY):     # every pod fits: filters agree
        work = [po.NodeMetrics(**{**n.__dict__}) for n in nodes]
        idxs, _ = po.greedy_place(work, pods, "resource_balanced", mode)
        assert [work[i].name for i in idxs] == g["binds"]
        for n, f in zip(work, g["final_table"]):
            assert n.pod_count == f["pod_count"] and bits(n.cpu_usage_percent) == f["cu_bits"]
            assert bits(n.available_memory) == f["available_memory_bits"]


def test_c_oracle_matches_py_oracle_capacity_filter():
    """The extension filter has no reference behaviour to pin; the two restatements of
    its spec must at least agree with each other bit-for-bit."""
    rng = np.random.Generator(np.random.PCG64(99))
    shapes = [(4000, 16, 30), (16000, 64, 58), (32000, 128, 110), (96000, 768, 250)]
    for trial in range(6):
        n, p = int(rng.integers(5, 60)), 80
        sh = rng.integers(0, 4, n)
        mp = np.array([shapes[k][2] for k in sh], dtype=np.int64)
        mp[rng.random(n) < 0.05] = 0
        pc = (rng.random(n) * (mp + 1)).astype(np.int64)
        arrays = dict(alloc_cpu_milli=np.array([shapes[k][0] for k in sh]),
                      alloc_mem_bytes=np.array([shapes[k][1] << 30 for k in sh]),
                      pod_count=pc, max_pods=mp, ready=(rng.random(n) > 0.1).astype(np.uint8),
                      cpu_usage_pct=np.array([float(po.usage_percent(int(a), int(b))) for a, b in zip(pc, mp)]),
                      mem_usage_pct=np.array([float(po.usage_percent(int(a), int(b))) for a, b in zip(pc, mp)]))
        rc = rng.integers(50, 120000, p)
        rm = rng.integers(64, 900 * 1024, p) << 20
        sid = rng.integers(0, 3, p)
        nodes = po.nodes_from_arrays(**arrays)
        pods = po.pods_from_arrays(rc, rm)
        names = [po.strategy_name(s) for s in sid]
        for mode in (0, 1):
            pi, ps = po.snapshot_place(nodes, pods, names, mode)
            ci, cs = c_oracle.snapshot(arrays, rc, rm, strategy_ids=sid, filter_mode=mode, threads=3)
            assert ci.tolist() == pi
            assert [bits(x) for x in cs] == [bits(x) for x in ps]
            work = po.nodes_from_arrays(**arrays)
            gi, gs = po.greedy_place(work, pods, names, mode)
            ci, cs, fpc, fcu, _ = c_oracle.greedy(arrays, rc, rm, strategy_ids=sid, filter_mode=mode)
            assert ci.tolist() == gi
            assert [bits(x) for x in cs] == [bits(x) for x in gs]
            assert fpc.tolist() == [w.pod_count for w in work]
            assert [bits(x) for x in fcu] == [bits(w.cpu_usage_percent) for w in work]


def test_oracle_restarted_from_a_checkpoint_continues_identically():
    """The checkpoint helper the full-size GPU greedy tests rely on (tests/helpers.py: state_after), pinned on
    the oracle itself: a run restarted from the state implied by its own first k decisions repeats the rest."""
    from k8s_llm_scheduler_b200 import synth
    from tests.helpers import check_greedy_from_checkpoint, oracle_greedy
    nodes, pods = synth.config3(300, 30_000)                       # fills up around pod 16 000
    o = oracle_greedy(nodes, pods, 0, 1)
    assert (o[0] < 0).any()
    for k in (0, 9_000, 15_500, 25_000):
        check_greedy_from_checkpoint(nodes, pods, o[0], o[1], k, 3_000, 0, 1)
    nodes, pods = synth.config5(1500, 40_000)
    o = oracle_greedy(nodes, pods, 0, 1, per_pod=True)
    for k in (0, 20_000, 37_000):
        check_greedy_from_checkpoint(nodes, pods, o[0], o[1], k, 3_000, 0, 1, per_pod=True)
