"""kp_snapshot_place_pruned: the first-feasible probe in front of the table scan must not change a single answer.
Compared with the CPU oracle (node index exact, score bit for bit) on the synthetic BASELINE shapes -- where every
feasible pod fits the best node and nothing is left to scan -- and on tables built so that the first fit lies
anywhere: inside the probe, behind it, or nowhere."""
import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200._lib import FILTER_CAPACITY, FILTER_READY
from k8s_llm_scheduler_b200.tables import NodeTable, PodTable
from tests.helpers import assert_same_f64, assert_same_idx, oracle_snapshot

pytestmark = pytest.mark.gpu


def pruned(nodes, pods, strategy, depth, per_pod=False, filter_mode=FILTER_CAPACITY):
    import torch
    from k8s_llm_scheduler_b200 import placement
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    strat = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else strategy
    idx, score, scanned = placement.snapshot_place_pruned(dn, rc, rm, strat, filter_mode, depth, scanned=True)
    return idx.cpu().numpy(), score.cpu().numpy(), scanned


def scattered_cluster(n, p, seed, tight=0.5):
    """Score and free capacity independent of each other: the best-scoring nodes are as likely to be full as any,
    so the first node a pod fits lies anywhere in the best-first order (or nowhere)."""
    rng = np.random.default_rng(seed)
    nodes = NodeTable.from_columns(
        alloc_cpu_milli=rng.integers(500, 64_000, n), alloc_mem_bytes=rng.integers(1 << 28, 256 << 30, n),
        pod_count=rng.integers(0, 110, n), max_pods=rng.choice([0, 30, 110, 110, 250], n),
        ready=(rng.random(n) > 0.03).astype(np.uint8), cpu_usage_pct=np.round(rng.random(n) * 100, 2),
        mem_usage_pct=np.round(rng.random(n) * 100, 2))
    hi_c, hi_m = int(64_000 * tight), int((256 << 30) * tight)
    pods = PodTable.from_columns(rng.integers(1, hi_c, p), rng.integers(1, hi_m, p), rng.integers(0, 4, p))
    return nodes, pods


@pytest.mark.parametrize("n,p", [(3, 50), (40, 3000), (700, 20_000), (5000, 30_000), (9000, 30_000), (20_000, 40_000)])
@pytest.mark.parametrize("depth", [1, 8, 64])
def test_first_fit_anywhere(n, p, depth):
    nodes, pods = scattered_cluster(n, p, seed=n + depth)
    for sid, per_pod in ((0, False), (1, False), (2, False), (0, True)):
        gi, gs, scanned = pruned(nodes, pods, sid, depth, per_pod)
        oi, os_ = oracle_snapshot(nodes, pods, sid, FILTER_CAPACITY, per_pod=per_pod)
        assert_same_idx(gi, oi, f"n={n} p={p} depth={depth} strategy={sid} per_pod={per_pod}")
        assert_same_f64(gs, os_)
        assert 0 <= sum(scanned) <= p


def test_probe_depths_split_the_work_differently_but_not_the_answers():
    nodes, pods = scattered_cluster(6000, 50_000, seed=77, tight=0.9)
    oi, os_ = oracle_snapshot(nodes, pods, 0, FILTER_CAPACITY)
    left = []
    for depth in (0, 1, 4, 16, 64, 1000):                         # 0: no probe (nothing is grouped); > 64 is clamped
        gi, gs, scanned = pruned(nodes, pods, 0, depth)
        assert_same_idx(gi, oi, f"depth {depth}")
        assert_same_f64(gs, os_)
        left.append(sum(scanned))
    assert left[0] == 0                                            # ungrouped call: the counters are not used
    assert left[1] >= left[2] >= left[3] >= left[4] == left[5] and left[1] > left[4]   # deeper probes leave fewer pods for the scan


def test_nothing_fits_the_probed_nodes():
    """The best-scoring nodes have no room at all: every pod goes on to the table scan."""
    n, p = 3000, 20_000
    rng = np.random.default_rng(5)
    usage = np.sort(rng.random(n) * 90)                            # node 0 scores best under least_loaded
    alloc = np.where(np.arange(n) < 200, 100, 32_000)              # ... and the first 200 have 0.1 cores
    nodes = NodeTable.from_columns(alloc_cpu_milli=alloc, alloc_mem_bytes=[64 << 30] * n, pod_count=[5] * n, max_pods=[110] * n,
                                   ready=[1] * n, cpu_usage_pct=usage, mem_usage_pct=usage)
    pods = PodTable.from_columns(rng.integers(500, 2000, p), rng.integers(1 << 20, 1 << 30, p))
    gi, gs, scanned = pruned(nodes, pods, 1, 64)
    oi, os_ = oracle_snapshot(nodes, pods, 1, FILTER_CAPACITY)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)
    assert scanned == [0, p, 0] and (gi >= 200).all()


def test_baseline_shapes_leave_nothing_to_scan():
    nodes, pods = synth.config5(5000, 60_000)
    gi, gs, scanned = pruned(nodes, pods, 0, 32, per_pod=True)
    oi, os_ = oracle_snapshot(nodes, pods, 0, FILTER_CAPACITY, per_pod=True)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)
    assert sum(scanned) == 0                                       # infeasible pods are rejected, the rest fit the best node
    nodes, pods = synth.config2()
    gi, gs, scanned = pruned(nodes, pods, 1, 32)
    oi, os_ = oracle_snapshot(nodes, pods, 1, FILTER_CAPACITY)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)


def test_ready_filter_ignores_the_probe_and_empty_inputs_work():
    nodes, pods = scattered_cluster(500, 4000, seed=3)
    gi, gs, _ = pruned(nodes, pods, 0, 32, filter_mode=FILTER_READY)
    oi, os_ = oracle_snapshot(nodes, pods, 0, FILTER_READY)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)
    none = PodTable.from_columns([], [])
    gi, gs, _ = pruned(nodes, none, 0, 32)
    assert len(gi) == 0
    dead = NodeTable.from_columns(alloc_cpu_milli=[8000] * 4, alloc_mem_bytes=[16 << 30] * 4, pod_count=[1] * 4, max_pods=[110] * 4,
                                  ready=[0] * 4, cpu_usage_pct=[1.0] * 4, mem_usage_pct=[1.0] * 4)
    gi, gs, _ = pruned(dead, pods, 0, 32)
    assert (gi == -1).all() and (gs == -1.0).all()


def test_plan_left_by_a_pruned_call_serves_place_again_and_the_session_option():
    import torch
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = scattered_cluster(4000, 30_000, seed=11)
    more = scattered_cluster(4000, 9000, seed=12)[1]
    pods, more = (PodTable.from_columns(t.req_cpu_milli, t.req_mem_bytes) for t in (pods, more))     # one strategy for all pods
    dn = placement.nodes_to_device(nodes, "cuda:0")
    ws = torch.empty(placement.snapshot_workspace_bytes(len(nodes), len(pods), True), dtype=torch.uint8, device="cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    gi, gs = placement.snapshot_place_pruned(dn, rc, rm, 0, FILTER_CAPACITY, 16, workspace=ws)
    oi, os_ = oracle_snapshot(nodes, pods, 0, FILTER_CAPACITY)
    assert_same_idx(gi.cpu().numpy(), oi)
    rc2, rm2 = torch.as_tensor(more.req_cpu_milli).cuda(), torch.as_tensor(more.req_mem_bytes).cuda()
    gi, gs = placement.snapshot_place_again(dn, rc2, rm2, 0, ws)
    oi, os_ = oracle_snapshot(nodes, more, 0, FILTER_CAPACITY)
    assert_same_idx(gi.cpu().numpy(), oi)
    assert_same_f64(gs.cpu().numpy(), os_)
    with placement.Session(max_nodes=4096, max_pods=1 << 16) as sess:
        sess.set_prune_depth(16)
        for sid in (0, 2):
            hi, hs = sess.snapshot(nodes, pods, sid, FILTER_CAPACITY)
            oi, os_ = oracle_snapshot(nodes, pods, sid, FILTER_CAPACITY)
            assert_same_idx(hi, oi)
            assert_same_f64(hs, os_)
        mixed = scattered_cluster(4000, 9000, seed=13)[1]            # per-pod strategy ids through the session, probed too
        hi, hs = sess.snapshot(nodes, mixed, 0, FILTER_CAPACITY)
        oi, os_ = oracle_snapshot(nodes, mixed, 0, FILTER_CAPACITY, per_pod=True)
        assert_same_idx(hi, oi)
        assert_same_f64(hs, os_)
        sess.set_nodes(nodes)                                       # resident table: the second call reuses the plan
        for _ in range(2):
            hi, hs = sess.snapshot(None, more, 0, FILTER_CAPACITY)
            oi, os_ = oracle_snapshot(nodes, more, 0, FILTER_CAPACITY)
            assert_same_idx(hi, oi)
            assert_same_f64(hs, os_)


@pytest.mark.parametrize("n,pinned", [(3000, True), (9000, False)])
def test_session_probe_with_pods_uploaded_in_pieces(n, pinned):
    """Two million pods through the session with the probe on: the columns go up in equal pieces, every piece is
    grouped / probed / scanned against the first piece's plan, answers written as the pieces finish."""
    import torch
    from k8s_llm_scheduler_b200.placement import Session
    from tests.helpers import device_snapshot
    p = 2_100_003
    nodes, pods = scattered_cluster(n, p, seed=90 + n, tight=0.9)
    uniform = PodTable.from_columns(pods.req_cpu_milli, pods.req_mem_bytes)
    want = device_snapshot(nodes, uniform, 0, FILTER_CAPACITY)               # the full scan (oracle-checked below on a sample)
    head = PodTable.from_columns(pods.req_cpu_milli[:30_000], pods.req_mem_bytes[:30_000])
    oi, _ = oracle_snapshot(nodes, head, 0, FILTER_CAPACITY)
    assert_same_idx(want[0][:30_000], oi)
    with Session(0, max_nodes=n, max_pods=p) as s:
        s.set_prune_depth(8)
        if pinned:
            hp = (torch.as_tensor(pods.req_cpu_milli).pin_memory(), torch.as_tensor(pods.req_mem_bytes).pin_memory(), None)
            out = (torch.empty(p, dtype=torch.int64).pin_memory(), torch.empty(p, dtype=torch.float64).pin_memory())
            for _ in range(2):
                out[0].fill_(-7)
                s.snapshot(nodes, hp, 0, FILTER_CAPACITY, out=out)
                assert_same_idx(out[0].numpy(), want[0])
                assert_same_f64(out[1].numpy(), want[1])
        else:
            gi, gs = s.snapshot(nodes, uniform, 0, FILTER_CAPACITY)
            assert_same_idx(gi, want[0])
            assert_same_f64(gs, want[1])
