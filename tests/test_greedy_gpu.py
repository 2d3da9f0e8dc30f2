"""Parity of the CUDA greedy-commit path (kp_greedy_place / kp_session_greedy) against the
reference's own bind sequences (tests/golden/greedy.json) and the CPU oracle."""
import json

import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200.tables import NodeTable, PodTable, strategy_id
from tests.helpers import (COND, assert_same_f64, assert_same_idx, bits, check_greedy_from_checkpoint,
                           device_greedy, oracle_greedy, state_after)

pytestmark = pytest.mark.gpu


def load(golden_dir, name):
    return json.loads((golden_dir / name).read_text())


def test_reference_bind_sequences(golden_dir):
    """The reference's own loop (real get_node_metrics + _fallback_decision + bind)."""
    for case in load(golden_dir, "greedy.json"):
        rows = [dict(name=ns["name"], cpu=ns["cpu"], memory=ns["memory"], pods=ns["pods"],
                     pod_count=ns["n_running"], conditions=COND[ns.get("cond", "ready")])
                for ns in case["nodes"]]
        nodes = NodeTable.from_allocatable(rows)
        p = case["n_pods"]
        pods = PodTable.from_columns([100] * p, [1 << 20] * p)
        idx, score, pc, cu, mu = device_greedy(nodes, pods, strategy_id(case["strategy"]), 0)
        assert idx.tolist() == case["idx_seq"], case["tag"]
        assert [bits(s) for s in score] == case["score_bits_seq"], case["tag"]
        assert pc.tolist() == case["final_pod_count"], case["tag"]


def test_config1_bind_sequence(golden_dir):
    g = load(golden_dir, "config1_trace.json")
    nodes, pods = synth.config1()
    for mode in (0, 1):
        idx, score, pc, cu, mu = device_greedy(nodes, pods, 0, mode)
        assert [nodes.names[i] for i in idx] == g["binds"] == ["minikube-m02", "minikube-m03", "minikube-m02"]
        assert pc.tolist() == [f["pod_count"] for f in g["final_table"]]
        assert [bits(x) for x in cu] == [f["cu_bits"] for f in g["final_table"]]


@pytest.mark.parametrize("n,p", [(1, 40), (3, 100), (60, 3000), (1024, 1500), (1025, 1100), (2500, 2100),
                                 (4096, 900), (5000, 2500), (6144, 600), (7000, 300)])
def test_random_clusters_against_oracle(n, p):
    nodes, pods = synth.config5(n, p, seed=2000 + n)
    for mode in (0, 1):
        for sid in (0, 1, 2):
            g = device_greedy(nodes, pods, sid, mode)
            o = oracle_greedy(nodes, pods, sid, mode)
            assert_same_idx(g[0], o[0], f"n={n} mode={mode} strategy={sid}")
            assert_same_f64(g[1], o[1])
            assert np.array_equal(g[2], o[2])
            assert_same_f64(g[3], o[3], "cpu usage")
            assert_same_f64(g[4], o[4], "mem usage")
        g = device_greedy(nodes, pods, 0, mode, per_pod=True)
        o = oracle_greedy(nodes, pods, 0, mode, per_pod=True)
        assert_same_idx(g[0], o[0], f"n={n} mode={mode} per-pod strategies")
        assert_same_f64(g[1], o[1])
        assert np.array_equal(g[2], o[2])


def test_config3_reduced_depletion():
    """Config-3 shape (homogeneous, pod_count ~ U{0..55}) with enough pods to fill nodes up."""
    nodes, pods = synth.config3(300, 30_000)
    g = device_greedy(nodes, pods, 0, 1)
    o = oracle_greedy(nodes, pods, 0, 1)
    assert_same_idx(g[0], o[0])
    assert_same_f64(g[1], o[1])
    assert np.array_equal(g[2], o[2])
    assert (g[0] < 0).any() and (g[2] <= 110).all()            # cluster fills, then pods fail


def test_config3_full_nodes_sampled_pods():
    nodes, pods = synth.config3(5000, 40_000)
    g = device_greedy(nodes, pods, 0, 1)
    o = oracle_greedy(nodes, pods, 0, 1)
    assert_same_idx(g[0], o[0])
    assert_same_f64(g[1], o[1])
    assert np.array_equal(g[2], o[2])


def test_round_robin_piles_on_one_node():
    nodes, pods = synth.config2(200, 500)
    idx, score, pc, *_ = device_greedy(nodes, pods, 2, 0)
    first = int(np.argmax(nodes.pod_count))
    assert (idx == first).all() and pc[first] == nodes.pod_count[first] + 500


def test_edge_cases():
    nodes, pods = synth.config2(20, 10)
    empty = PodTable.from_columns([], [])
    idx, score, pc, cu, mu = device_greedy(nodes, empty, 0, 1)
    assert idx.size == 0 and np.array_equal(pc, nodes.pod_count)
    nodes.ready[:] = 0
    idx, score, pc, *_ = device_greedy(nodes, pods, 0, 1)
    assert (idx == -1).all() and (score == -1.0).all() and np.array_equal(pc, nodes.pod_count)
    # custom usage that does not follow A3 is overwritten by the first commit
    nodes, pods = synth.config2(4, 6)
    nodes.cpu_usage_pct[:] = [10.0, 3.0, 3.0, 77.0]
    nodes.mem_usage_pct[:] = [1.0, 3.0, 2.5, 0.0]
    for sid in (0, 1):
        g = device_greedy(nodes, pods, sid, 0)
        o = oracle_greedy(nodes, pods, sid, 0)
        assert_same_idx(g[0], o[0])
        assert_same_f64(g[1], o[1])
        assert_same_f64(g[3], o[3])


def test_session_greedy_matches_device_api():
    from k8s_llm_scheduler_b200.placement import Session
    nodes, pods = synth.config5(1500, 4000, seed=9)
    with Session(0, max_nodes=2048, max_pods=4096) as s:
        h = s.greedy(nodes, pods, 0, 1)
        d = device_greedy(nodes, pods, 0, 1, per_pod=True)
        assert_same_idx(h[0], d[0])
        assert_same_f64(h[1], d[1])
        assert np.array_equal(h[2], d[2])
        # caller-owned (pinned) result buffers
        import torch
        out = tuple(torch.empty(k, dtype=t).pin_memory() for k, t in ((4000, torch.int64), (4000, torch.float64), (1500, torch.int64),
                                                                      (1500, torch.float64), (1500, torch.float64)))
        assert s.greedy(nodes, pods, 0, 1, out=out) is out
        assert_same_idx(out[0].numpy(), d[0])
        assert_same_f64(out[1].numpy(), d[1])
        assert np.array_equal(out[2].numpy(), d[2])
        assert_same_f64(out[3].numpy(), d[3])
        with pytest.raises(TypeError):
            s.greedy(nodes, pods, 0, 1, out=out[:4] + (np.empty(7),))


# ---- fast path (kp_greedy_fast.cu): integer state machine, clusters, fallbacks --------------
def _check_greedy(nodes, pods, sid, mode, per_pod=False, what=""):
    g = device_greedy(nodes, pods, sid, mode, per_pod=per_pod)
    o = oracle_greedy(nodes, pods, sid, mode, per_pod=per_pod)
    assert_same_idx(g[0], o[0], what)
    assert_same_f64(g[1], o[1], what + " score")
    assert np.array_equal(g[2], o[2]), what + " final pod counts"
    assert_same_f64(g[3], o[3], what + " cpu usage")
    assert_same_f64(g[4], o[4], what + " mem usage")


@pytest.mark.parametrize("cluster", ["1", "2", "4", "8"])
def test_fast_path_on_clusters(cluster, monkeypatch):
    """Force the thread-block-cluster variants (DSMEM exchange) on a table one CTA could hold."""
    monkeypatch.setenv("KP_GREEDY_CLUSTER", cluster)
    nodes, pods = synth.config5(3000, 6000, seed=31)
    for sid in (0, 1, 2):
        _check_greedy(nodes, pods, sid, 1, what=f"cluster={cluster} strategy={sid}")
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what=f"cluster={cluster} per-pod")


def test_fast_path_large_tables_and_depletion():
    nodes, pods = synth.config5(20_000, 30_000, seed=32)          # one CTA, 80 nodes per thread
    _check_greedy(nodes, pods, 0, 1, what="20k uniform")
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what="20k per-pod (cluster of 2)")
    nodes, pods = synth.config5(8_000, 6_000, seed=36)             # one CTA, 40 nodes per thread
    _check_greedy(nodes, pods, 1, 1, what="8k uniform")
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what="8k per-pod")
    nodes, pods = synth.config5(30_000, 5_000, seed=37)            # past one CTA's registers: cluster of 2 / of 4
    _check_greedy(nodes, pods, 0, 1, what="30k uniform (cluster of 2)")
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what="30k per-pod (cluster of 4)")
    nodes, pods = synth.config3(400, 50_000)                       # fills up: trajectories run to max_pods
    _check_greedy(nodes, pods, 0, 1, what="depletion")
    g = device_greedy(nodes, pods, 0, 1)
    assert (g[0] < 0).any() and (g[2] <= 110).all()


@pytest.mark.parametrize("cluster", ["1", "2"])
def test_dead_pod_bound_follows_the_live_states(cluster, monkeypatch):
    """The loop refreshes its "fits nothing" bound from the live node states at every pod batch.  It must stay exact
    (a) when the cluster fills up and the rest of the pods is answered without a scan, per-pod strategies included;
    (b) when a node's FIRST commit makes it roomier -- the caller's usage figures (99 %) are replaced by the
        reference's pod_count / max_pods formula (scheduler.py:150-151), so pods that fit no node now fit later;
    (c) when thresholds grow along a trajectory (negative allocatable): the refresh must not be used at all."""
    monkeypatch.setenv("KP_GREEDY_CLUSTER", cluster)
    rng = np.random.default_rng(8)
    nodes, pods = synth.config5(300, 40_000, seed=41)                                      # (a) ~15 k free slots, 40 k pods
    for sid in (0, 2):
        _check_greedy(nodes, pods, sid, 1, what=f"fills up, strategy {sid}")
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what="fills up, per-pod strategies")
    n = 64                                                                                 # (b)
    busy = NodeTable.from_columns(alloc_cpu_milli=[32_000] * n, alloc_mem_bytes=[128 << 30] * n, pod_count=[1] * n, max_pods=[110] * n,
                                  ready=[1] * n, cpu_usage_pct=[99.0] * n, mem_usage_pct=[99.0] * n)
    small = rng.random(6000) < 0.5
    small[:3] = False                                              # the first pods are large: no node has been reset yet
    mixed = PodTable.from_columns(np.where(small, 100, 8000), np.where(small, 64 << 20, 16 << 30))
    for sid in (0, 1, 2):
        _check_greedy(busy, mixed, sid, 1, what=f"first commit frees capacity, strategy {sid}")
    g = device_greedy(busy, mixed, 0, 1)
    assert (g[0][~small] >= 0).any() and (g[0][~small] < 0).any()      # large pods: rejected at first, placed once a node was reset
    odd = NodeTable.from_columns(alloc_cpu_milli=[-4000] * 8 + [16_000] * 40, alloc_mem_bytes=[64 << 30] * 48, pod_count=[2] * 48,    # (c)
                                 max_pods=[60] * 48, ready=[1] * 48, cpu_usage_pct=[10.0] * 48, mem_usage_pct=[10.0] * 48)
    neg = PodTable.from_columns(rng.integers(-4500, 3000, 5000), rng.integers(1, 1 << 30, 5000))
    for sid in (0, 1):
        _check_greedy(odd, neg, sid, 1, what=f"negative allocatable, strategy {sid}")


def test_fast_path_matches_fp64_kernel(monkeypatch):
    nodes, pods = synth.config5(2500, 5000, seed=33)
    fast = device_greedy(nodes, pods, 0, 1, per_pod=True)
    monkeypatch.setenv("KP_GREEDY_NO_FAST", "1")
    slow = device_greedy(nodes, pods, 0, 1, per_pod=True)
    for a, b in zip(fast, slow):
        assert np.array_equal(np.asarray(a).view(np.uint64) if a.dtype == np.float64 else a,
                              np.asarray(b).view(np.uint64) if b.dtype == np.float64 else b)


def test_fast_path_falls_back_when_preconditions_fail():
    rng = np.random.Generator(np.random.PCG64(34))
    n, p = 1500, 4000
    mp = rng.integers(20, 200, n)
    pc = (rng.random(n) * mp).astype(np.int64)
    u = synth.usage_a3(pc, mp)
    # every node its own shape (> 128 shapes): the device-side check sends the work to the fp64 kernel
    nodes = NodeTable.from_columns(alloc_cpu_milli=rng.integers(2000, 64000, n), alloc_mem_bytes=rng.integers(4, 512, n) << 30,
                                   pod_count=pc, max_pods=mp, ready=(rng.random(n) > 0.05).astype(np.uint8),
                                   cpu_usage_pct=u, mem_usage_pct=u.copy())
    pods = PodTable.from_columns(rng.integers(50, 8000, p), rng.integers(64, 32 * 1024, p) << 20, rng.integers(0, 3, p))
    _check_greedy(nodes, pods, 0, 1, what="many shapes")
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what="many shapes per-pod")
    # few shapes but usage that does not follow scheduler.py:150-151: initial states are their own entries
    nodes, pods = synth.config5(800, 5000, seed=35)
    nodes.cpu_usage_pct[:] = np.round(rng.random(800) * 60, 2)
    nodes.mem_usage_pct[:] = np.round(rng.random(800) * 60, 2)
    for sid in (0, 1):
        _check_greedy(nodes, pods, sid, 1, what=f"custom usage strategy={sid}")
    # max_pods too large for a trajectory table
    nodes, pods = synth.config2(50, 3000)
    nodes.max_pods[:] = 1 << 25
    _check_greedy(nodes, pods, 0, 1, what="huge max_pods")


@pytest.mark.parametrize("strategy", [0, 1])
def test_batched_prefix_and_hand_over_to_the_loop(strategy):
    """Exact batching (kp_greedy_fast.cu): while every pod fits every reachable node state the decisions are one
    sort; the first pod that does not ends the prefix and the sequential loop continues on the advanced states.
    The violating pod is placed at several depths (before the minimum run length, inside the run, after the
    cluster has filled up); every variant must equal the oracle pod for pod, state for state."""
    nodes, base = synth.config3(600, 40_000, seed=90 + strategy)           # 600 x ~82 free slots: fills around pod 49 000
    for where in (None, 100, 5_000, 22_222, 39_999):
        pods = PodTable.from_columns(base.req_cpu_milli.copy(), base.req_mem_bytes.copy())
        if where is not None:
            pods.req_cpu_milli[where] = 17_000                             # fits an empty node (32 cores x 0.5+), not a loaded one
            pods.req_cpu_milli[where + 1 if where + 1 < len(pods) else where] = 40_000   # fits nothing
        _check_greedy(nodes, pods, strategy, 1, what=f"violation at {where}")
    nodes, pods = synth.config3(300, 60_000, seed=93)                      # more pods than free slots: the entries run out
    _check_greedy(nodes, pods, strategy, 1, what="entries run out inside the prefix")
    # usage that does not follow scheduler.py:150-151: a node's score may RISE at its first commit -> no batching
    nodes, pods = synth.config3(500, 30_000, seed=94)
    rng = np.random.Generator(np.random.PCG64(95))
    nodes.cpu_usage_pct[:] = np.round(rng.random(500) * 90, 2)
    nodes.mem_usage_pct[:] = np.round(rng.random(500) * 90, 2)
    _check_greedy(nodes, pods, strategy, 1, what="custom usage")
    # heterogeneous shapes, all pods small enough for the tightest state of the smallest shape
    nodes, pods = synth.config5(3000, 50_000, seed=96, infeasible=0.0)
    pods.req_cpu_milli[:] = np.minimum(pods.req_cpu_milli, 1_900)
    pods.req_mem_bytes[:] = np.minimum(pods.req_mem_bytes, 7 << 30)
    _check_greedy(nodes, pods, strategy, 1, what="heterogeneous shapes")


def test_full_size_config3_greedy():
    """BASELINE config 3 at full size (5k nodes x 1M pods, resource_balanced, sequential commit):
    conservation properties at full size; exact parity with the oracle on the first 30k pods (a
    greedy prefix does not depend on the pods behind it)."""
    nodes, pods = synth.config3(5000, 1_000_000)
    idx, score, pc, cu, mu = device_greedy(nodes, pods, 0, 1)
    placed = idx >= 0
    per_node = np.bincount(idx[placed], minlength=len(nodes))
    assert np.array_equal(pc - nodes.pod_count, per_node)                 # every commit is accounted for
    assert (pc <= np.maximum(nodes.max_pods, nodes.pod_count)).all()      # nobody is filled past max_pods
    assert (score[~placed] == -1.0).all()
    touched = per_node > 0
    assert_same_f64(cu[touched], synth.usage_a3(pc[touched], nodes.max_pods[touched]), "final cpu usage")
    assert_same_f64(mu[~touched], nodes.mem_usage_pct[~touched], "untouched mem usage")
    again = device_greedy(nodes, pods, 0, 1)
    assert np.array_equal(again[0], idx)
    head = PodTable.from_columns(pods.req_cpu_milli[:30_000], pods.req_mem_bytes[:30_000])
    o = oracle_greedy(nodes, head, 0, 1)
    assert_same_idx(idx[:30_000], o[0])
    assert_same_f64(score[:30_000], o[1])
    # Checkpointed oracle runs: greedy is Markov in the node state, so the oracle restarted from the state
    # the device's own decisions imply must continue identically.  The cluster has 414 448 free pod slots:
    # checkpoints sit in the regime where trajectories run towards max_pods, straddle the depletion point
    # (the last placement), and in the all-rejected tail.
    free_slots = int((nodes.max_pods - nodes.pod_count)[nodes.ready != 0].clip(min=0).sum())
    last_placed = int(np.nonzero(placed)[0][-1])
    assert free_slots == 414_448 and int(placed.sum()) == free_slots and last_placed == free_slots - 1
    for k in (200_000, 400_000, last_placed - 2_500, 700_000, 995_000):
        check_greedy_from_checkpoint(nodes, pods, idx, score, k, 5_000, 0, 1, what=f"c3 checkpoint {k}")
    final = state_after(nodes, idx)
    assert np.array_equal(final.pod_count, pc)


def test_full_size_config5_greedy_form():
    """BASELINE config 5 in greedy form at full size: 20k heterogeneous nodes x 2M pods, per-pod strategies
    (cluster of 2 CTAs with the st.async exchange), 30 % infeasible pods.  Conservation properties over the
    whole run plus checkpointed oracle runs of 5 000 pods each, from the head to the tail."""
    nodes, pods = synth.config5(20_000, 2_000_000)
    idx, score, pc, cu, mu = device_greedy(nodes, pods, 0, 1, per_pod=True)
    placed = idx >= 0
    per_node = np.bincount(idx[placed], minlength=len(nodes))
    assert np.array_equal(pc - nodes.pod_count, per_node)
    assert (pc <= np.maximum(nodes.max_pods, nodes.pod_count)).all()
    assert (per_node[nodes.ready == 0] == 0).all()
    assert (score[~placed] == -1.0).all()
    touched = per_node > 0
    assert_same_f64(cu[touched], synth.usage_a3(pc[touched], nodes.max_pods[touched]), "final cpu usage")
    assert_same_f64(cu[~touched], nodes.cpu_usage_pct[~touched], "untouched cpu usage")
    # round_robin pods pile onto the fullest fitting node, the other two spread: all three kinds place pods
    for sid in (0, 1, 2):
        assert placed[pods.strategy_ids == sid].any()
    last_placed = int(np.nonzero(placed)[0][-1])
    cps = sorted({0, 250_000, 1_000_000, max(0, last_placed - 2_500), 1_600_000, 1_995_000})
    for k in cps:
        check_greedy_from_checkpoint(nodes, pods, idx, score, k, 5_000, 0, 1, per_pod=True, what=f"c5 checkpoint {k}")


def test_tables_beyond_the_fast_path_capacity_use_the_fp64_kernel():
    """More nodes than eight CTAs hold in registers (163 840; 81 920 with per-pod strategies)."""
    nodes, pods = synth.config5(170_000, 2100, seed=38)
    _check_greedy(nodes, pods, 0, 1, what="170k uniform")
    nodes, pods = synth.config5(90_000, 2100, seed=39)
    _check_greedy(nodes, pods, 0, 1, per_pod=True, what="90k per-pod")
    _check_greedy(nodes, pods, 1, 1, what="90k uniform (cluster of 8)")


def test_greedy_on_a_second_device_of_the_same_process(monkeypatch):
    """Opt-in shared-memory sizes are a per-device attribute: the second GPU a process uses must get them too
    (fp64 kernel with more than 48 KB of cold state, fast path with its trajectory tables)."""
    import torch
    from k8s_llm_scheduler_b200 import placement
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    nodes, pods = synth.config5(4000, 5000, seed=41)
    want = oracle_greedy(nodes, pods, 0, 1)
    want_pp = oracle_greedy(nodes, pods, 0, 1, per_pod=True)
    for no_fast in ("0", "1"):
        monkeypatch.setenv("KP_GREEDY_NO_FAST", no_fast)
        for dev in ("cuda:0", "cuda:1"):
            dn = placement.nodes_to_device(nodes, dev)
            rc, rm = torch.as_tensor(pods.req_cpu_milli).to(dev), torch.as_tensor(pods.req_mem_bytes).to(dev)
            with torch.cuda.device(dev):
                g = placement.greedy_place(dn, rc, rm, 0, 1)
                gp = placement.greedy_place(dn, rc, rm, torch.as_tensor(pods.strategy_ids).to(dev), 1)
                torch.cuda.synchronize()
            assert_same_idx(g[0].cpu().numpy(), want[0], f"{dev} no_fast={no_fast}")
            assert np.array_equal(g[2].cpu().numpy(), want[2])
            assert_same_idx(gp[0].cpu().numpy(), want_pp[0], f"{dev} no_fast={no_fast} per-pod")
