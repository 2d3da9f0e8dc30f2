"""Node-sharded snapshot placement (kp_exchange.cu): W ranks, each with a contiguous slice of the node list, the
per-pod winner reduced across ranks over peer memory.  Here the W ranks are emulated on ONE GPU -- every rank
has its own exchange buffer and runs its own kernels, the three steps driven in lock step -- and the assembled
result must be the oracle's answer on the whole list, bit for bit.  (Real multi-GPU runs: tools/dist_check.py
under torchrun, and `bench.py --workload c5node`.)"""
import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200.tables import NODE_COLUMNS, NodeTable, PodTable
from tests.helpers import assert_same_f64, assert_same_idx, oracle_snapshot

pytestmark = pytest.mark.gpu


def run_emulated(nodes: NodeTable, pods: PodTable, world: int, strategy, per_pod: bool, rounds: int = 1, cuts=None, prune=0):
    import torch
    from k8s_llm_scheduler_b200 import dist as kd, placement
    dev = torch.device("cuda:0")
    n, p = len(nodes), len(pods)
    cuts = cuts or [kd.shard_bounds(n, r, world) for r in range(world)]
    group = kd.NodeShardedPlacer.local_group([dev] * world, max_pods=p, max_local_nodes=max(1, max(b - a for a, b in cuts)))
    for g in group:
        g.set_prune_depth(prune if g.rank % 2 == 0 else 0)     # the probe is a per-rank choice: mix ranks with and without
    full = placement.nodes_to_device(nodes, dev)
    slices = [kd.slice_nodes(full, a, b) for a, b in cuts]
    rc, rm = torch.as_tensor(pods.req_cpu_milli).to(dev), torch.as_tensor(pods.req_mem_bytes).to(dev)
    strat = torch.as_tensor(pods.strategy_ids).to(dev) if per_pod else strategy
    for _ in range(rounds):                                    # several snapshots in a row: epochs, alternating buffers
        for g, (a, _) in zip(group, cuts):
            g.keys(slices[g.rank], a, rc, rm, strat, 1)
        for g in group:
            g.scan(slices[g.rank], rc, rm, strat, 1)
        parts = [g.reduce() for g in group]
    torch.cuda.synchronize()
    for g in group:
        g.check()
    assert [(lo, hi) for lo, hi, _, _ in parts] == [kd.shard_bounds(p, r, world) for r in range(world)]
    idx = np.concatenate([i.cpu().numpy() for _, _, i, _ in parts])
    score = np.concatenate([s.cpu().numpy() for _, _, _, s in parts])
    for g in group:
        g.close()
    return idx, score


@pytest.mark.parametrize("n,p,world", [(2500, 20_001, 1), (2500, 20_001, 2), (1000, 5_000, 3), (20_000, 30_000, 8),
                                       (20_000, 12_345, 2), (70, 999, 8)])
def test_emulated_ranks_match_the_oracle(n, p, world):
    nodes, pods = synth.config5(n, p, seed=600 + n + world)
    for sid in (0, 1, 2):
        gi, gs = run_emulated(nodes, pods, world, sid, False)
        oi, os_ = oracle_snapshot(nodes, pods, sid, 1)
        assert_same_idx(gi, oi, f"n={n} world={world} strategy={sid}")
        assert_same_f64(gs, os_)
    gi, gs = run_emulated(nodes, pods, world, 0, True, rounds=3)
    oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
    assert_same_idx(gi, oi, f"n={n} world={world} per-pod strategies")
    assert_same_f64(gs, os_)


@pytest.mark.parametrize("n,p,world", [(2500, 20_001, 2), (20_000, 30_000, 8), (9000, 40_000, 3)])
def test_emulated_ranks_with_the_first_feasible_probe(n, p, world):
    """kp_exchange_set_prune_depth: ranks that find their local winner by the probe and ranks that scan, same answer.
    Score and capacity independent, so the local first fit lies anywhere in a slice."""
    rng = np.random.default_rng(n + world)
    nodes = NodeTable.from_columns(
        alloc_cpu_milli=rng.integers(500, 64_000, n), alloc_mem_bytes=rng.integers(1 << 28, 256 << 30, n),
        pod_count=rng.integers(0, 110, n), max_pods=rng.choice([0, 30, 110, 250], n), ready=(rng.random(n) > 0.03).astype(np.uint8),
        cpu_usage_pct=np.round(rng.random(n) * 100, 2), mem_usage_pct=np.round(rng.random(n) * 100, 2))
    pods = PodTable.from_columns(rng.integers(1, 60_000, p), rng.integers(1, 240 << 30, p), rng.integers(0, 4, p))
    for sid, per_pod in ((0, False), (2, False), (0, True)):
        gi, gs = run_emulated(nodes, pods, world, sid, per_pod, rounds=2, prune=8)
        oi, os_ = oracle_snapshot(nodes, pods, sid, 1, per_pod=per_pod)
        assert_same_idx(gi, oi, f"n={n} world={world} strategy={sid} per_pod={per_pod}")
        assert_same_f64(gs, os_)


def test_ties_across_ranks_go_to_the_first_node_in_list_order():
    """All nodes equal: every rank's local winner has the same score; the global winner is the first eligible
    node of the whole list, wherever the cuts fall (scheduler.py:547)."""
    nodes = synth.all_equal_nodes(4000)
    nodes.ready[:1300] = 0                                     # the whole first slice of a 4-way split and part of the second
    _, pods = synth.config2(10, 3000)
    for world in (2, 4, 7):
        for sid in (0, 1, 2):
            gi, _ = run_emulated(nodes, pods, world, sid, False)
            assert (gi == 1300).all(), (world, sid)


def test_uneven_and_empty_slices():
    nodes, pods = synth.config5(900, 4000, seed=77)
    want = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
    for cuts in ([(0, 0), (0, 900)], [(0, 1), (1, 1), (1, 650), (650, 900)], [(0, 899), (899, 900), (900, 900)]):
        gi, gs = run_emulated(nodes, pods, len(cuts), 0, True, cuts=cuts)
        assert_same_idx(gi, want[0], str(cuts))
        assert_same_f64(gs, want[1])
    # more ranks than pods: some shards are empty
    few = PodTable.from_columns(pods.req_cpu_milli[:3], pods.req_mem_bytes[:3])
    gi, _ = run_emulated(nodes, few, 5, 1, False)
    assert_same_idx(gi, oracle_snapshot(nodes, few, 1, 1)[0])


def test_two_real_gpus_in_one_process():
    """The same exchange with each rank on its own device (peer access inside one process): kernels of different
    ranks now really run at the same time and wait for each other's flags."""
    import torch
    from k8s_llm_scheduler_b200 import dist as kd, placement
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    nodes, pods = synth.config5(6000, 80_000, seed=88)
    n, p = len(nodes), len(pods)
    devs = [torch.device("cuda:0"), torch.device("cuda:1")]
    cuts = [kd.shard_bounds(n, r, 2) for r in range(2)]
    group = kd.NodeShardedPlacer.local_group(devs, max_pods=p, max_local_nodes=max(b - a for a, b in cuts))
    # every buffer exists before the first kernel that waits for a peer is launched (an allocation may have to
    # wait for running kernels)
    args, outs = [], []
    for g, dev, (a, b) in zip(group, devs, cuts):
        full = placement.nodes_to_device(nodes, dev)
        args.append((kd.slice_nodes(full, a, b), a, torch.as_tensor(pods.req_cpu_milli).to(dev),
                     torch.as_tensor(pods.req_mem_bytes).to(dev), torch.as_tensor(pods.strategy_ids).to(dev)))
        lo, hi = kd.shard_bounds(p, g.rank, 2)
        outs.append((torch.empty(hi - lo, dtype=torch.int64, device=dev), torch.empty(hi - lo, dtype=torch.float64, device=dev)))
        with torch.cuda.device(dev):
            g._structs(args[-1][0], args[-1][2], args[-1][3], args[-1][4])        # workspace allocated up front
    for dev in devs:
        torch.cuda.synchronize(dev)
    parts = []
    for rounds in range(3):
        parts = []
        for g, dev, arg, out in zip(group, devs, args, outs):
            with torch.cuda.device(dev):
                parts.append(g.place(*arg, 1, out=out))
        for dev in devs:
            torch.cuda.synchronize(dev)
    for g in group:
        g.check()
    idx = np.concatenate([q[2].cpu().numpy() for q in parts])
    score = np.concatenate([q[3].cpu().numpy() for q in parts])
    oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
    assert_same_idx(idx, oi)
    assert_same_f64(score, os_)
