"""bench.py's bookkeeping (what is counted, how work is split): CPU only, no kernels."""
import importlib.util
from pathlib import Path

import numpy as np

from k8s_llm_scheduler_b200 import synth

_spec = importlib.util.spec_from_file_location("bench", Path(__file__).resolve().parent.parent / "bench.py")
bench = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(bench)


def test_shards_cover_the_pods_once():
    for n, w in ((10, 3), (10_000_000, 8), (7, 8), (0, 4)):
        spans = [bench.shard_bounds(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1


def test_local_pods_are_contiguous_shards_of_the_same_job():
    nodes, pods = synth.config5(50, 1001)
    meta = dict(mode="snapshot", per_pod=True, filter=1, strategy=0)
    parts = [bench.local_pods(pods, meta, r, 4) for r in range(4)]
    assert np.array_equal(np.concatenate([p.req_cpu_milli for p, _ in parts]), pods.req_cpu_milli)
    assert np.array_equal(np.concatenate([p.strategy_ids for p, _ in parts]), pods.strategy_ids)
    assert [lo for _, lo in parts] == [bench.shard_bounds(1001, r, 4)[0] for r in range(4)]
    whole, lo = bench.local_pods(pods, dict(meta, mode="greedy"), 2, 4)           # greedy: replicas
    assert whole is pods and lo == 0
    whole, lo = bench.local_pods(pods, dict(meta, node_sharded=True), 2, 4)       # node-sharded: every rank sees all pods
    assert whole is pods


def test_eligible_nodes_and_scanned_pods_follow_the_kernels_rules():
    nodes, pods = synth.config5(4000, 20_000)
    meta = dict(mode="snapshot", per_pod=True, filter=1, strategy=0)
    el = bench.eligible_nodes(nodes, meta)
    assert el == int(((nodes.ready != 0) & (nodes.pod_count < nodes.max_pods)).sum()) < len(nodes)
    assert bench.eligible_nodes(nodes, dict(meta, filter=0)) == int((nodes.ready != 0).sum())
    scanned = bench.pods_scanned(nodes, pods, meta)
    # the oversized pods of config 5 (about 30 %) are larger than the largest node: rejected without a scan
    assert 0.65 * len(pods) < scanned < 0.75 * len(pods)
    assert bench.pods_scanned(nodes, pods, dict(meta, per_pod=False)) == len(pods)     # one strategy: every pod is scanned
    from oracle import c_oracle
    idx, _ = c_oracle.snapshot(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes, strategy_ids=pods.strategy_ids,
                               filter_mode=1, threads=4)
    assert (idx >= 0).sum() <= scanned                                                  # nothing that is placed was "rejected early"


def test_config_block_is_the_same_in_both_arms():
    nodes, pods, meta = bench.make_workload("c2")
    a = bench.config_block(meta, nodes, pods, 1)
    b = bench.config_block(meta, nodes, pods, 8)
    assert a == b and a["nodes"] == 1000 and a["pods"] == 100_000 and a["baseline_config"] == "configs[1]"
    assert bench.DEFAULT_WORKLOAD == "c4" and bench.scaling_of(meta) == "strong"
