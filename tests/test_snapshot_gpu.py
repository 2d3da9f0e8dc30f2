"""Parity of the CUDA snapshot path (kp_snapshot_place / kp_node_scores / kp_session_snapshot)
against the reference-generated golden vectors and the CPU oracle.  Node indices must match
exactly; scores are compared BIT-FOR-BIT (the north star allows 1e-9 relative)."""
import ctypes
import json

import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200.tables import NodeTable, PodTable, strategy_id
from tests.helpers import (assert_same_f64, assert_same_idx, bits, device_snapshot, from_bits,
                           oracle_snapshot)

pytestmark = pytest.mark.gpu
STRATS = ["resource_balanced", "least_loaded", "round_robin", "no_such_strategy"]


def load(golden_dir, name):
    return json.loads((golden_dir / name).read_text())


def table_from_specs(specs):
    """specs: list of (pc, mp, cu, mu, ready)."""
    n = len(specs)
    return NodeTable.from_columns(
        alloc_cpu_milli=[8000] * n, alloc_mem_bytes=[16374576 * 1024] * n,
        pod_count=[s[0] for s in specs], max_pods=[s[1] for s in specs],
        ready=[s[4] for s in specs], cpu_usage_pct=[s[2] for s in specs],
        mem_usage_pct=[s[3] for s in specs])


def test_node_scores_match_reference_bits(golden_dir):
    import torch
    from k8s_llm_scheduler_b200 import placement
    g = load(golden_dir, "kat_scores.json")
    rows = g["a3_usage"]
    specs = [(r["pc"], r["mp"], from_bits(r["usage_bits"]), from_bits(r["usage_bits"]), 1) for r in rows]
    specs += [(r["pc"], r["mp"], r["cu"], r["mu"], 1) for r in g["custom_usage"]]
    allrows = rows + g["custom_usage"]
    dn = placement.nodes_to_device(table_from_specs(specs), "cuda:0")
    for sid, s in enumerate(STRATS[:3]):
        score, _, _, elig = placement.node_scores(dn, sid, filter_mode=0)
        score, elig = score.cpu().numpy(), elig.cpu().numpy()
        for i, r in enumerate(allrows):
            assert bool(elig[i]) == r[s]["selected"], (i, s)
            if r[s]["selected"]:
                assert bits(score[i]) == r[s]["score_bits"], (r["pc"], r["mp"], s)


def test_reference_decisions_ready_filter(golden_dir):
    """Every golden decision of the unmodified reference, through the GPU path."""
    for case in load(golden_dir, "decisions.json"):
        if not case["nodes"]:
            continue
        specs = [(n["pc"], n["mp"], from_bits(n["cu_bits"]), from_bits(n["mu_bits"]),
                  int(n["cond"] in ("ready", "ready_second"))) for n in case["nodes"]]
        nodes = table_from_specs(specs)
        pods = PodTable.from_columns([100, 250000, 1], [1 << 20, 1 << 45, 0])
        for s in STRATS:
            exp = case["expect"][s]
            idx, score = device_snapshot(nodes, pods, strategy_id(s), filter_mode=0)
            assert idx.tolist() == [exp["idx"]] * 3, (case["tag"], s)
            assert [bits(x) for x in score] == [exp["score_bits"]] * 3, (case["tag"], s)


@pytest.mark.parametrize("n,p", [(1, 1), (3, 5), (7, 257), (100, 3000), (1000, 2049), (4096, 700),
                                 (4097, 1500), (5000, 300), (9001, 4099), (8192, 600), (8193, 600),
                                 (20000, 1000),
                                 # register-slice boundaries of place_small_kernel (56 words x 32 lanes per slice, word
                                 # counts in steps of 8) with pod counts off every group / pass size
                                 (255, 31), (256, 33), (257, 1001), (1792, 1001), (1793, 33), (2048, 4100), (3585, 31),
                                 (7169, 129), (8191, 3)])
def test_random_clusters_capacity_filter(n, p):
    nodes, pods = synth.config5(n, p, seed=1000 + n)
    for sid in (0, 1, 2):
        gi, gs = device_snapshot(nodes, pods, sid, 1)
        oi, os_ = oracle_snapshot(nodes, pods, sid, 1)
        assert_same_idx(gi, oi, f"n={n} p={p} strategy={sid}")
        assert_same_f64(gs, os_)
    gi, gs = device_snapshot(nodes, pods, 0, 1, per_pod=True)
    oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
    assert_same_idx(gi, oi, f"n={n} p={p} per-pod strategies")
    assert_same_f64(gs, os_)
    gi, gs = device_snapshot(nodes, pods, 0, 0, per_pod=True)      # reference filter, mixed
    oi, os_ = oracle_snapshot(nodes, pods, 0, 0, per_pod=True)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)


def test_config1_snapshot():
    nodes, pods = synth.config1()
    for mode in (0, 1):
        idx, score = device_snapshot(nodes, pods, 0, mode)
        assert idx.tolist() == [1, 1, 1]                      # m02 (ties with m03, first wins)
        assert [bits(s) for s in score] == ["3fef9f2f79f2f79f"] * 3


def test_config2_full_size_least_loaded():
    """BASELINE config 2 at full size: 1k nodes x 100k pods, every pod checked."""
    nodes, pods = synth.config2()
    gi, gs = device_snapshot(nodes, pods, 1, 1)
    oi, os_ = oracle_snapshot(nodes, pods, 1, 1)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)
    assert (gi >= 0).all()


def test_config5_reduced_mixed_strategies():
    nodes, pods = synth.config5(5000, 60_000)
    gi, gs = device_snapshot(nodes, pods, 0, 1, per_pod=True)
    oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
    assert_same_idx(gi, oi)
    assert_same_f64(gs, os_)
    assert 0.27 < (gi < 0).mean() < 0.33


@pytest.mark.parametrize("n,p", [(2500, 600_000), (8192, 270_000), (1025, 300_001), (3000, 262_144), (900, 530_000)])
def test_small_tables_with_many_pods(n, p):
    """One rank's share of a node-sharded cluster: a small table under hundreds of thousands of pods, uniform strategy
    (both placement kernels, either side of 2^19 pods) and per-pod strategies."""
    nodes, pods = synth.config5(n, p, seed=77 + n)
    for sid, per_pod in ((0, False), (1, False), (0, True)):
        gi, gs = device_snapshot(nodes, pods, sid, 1, per_pod=per_pod)
        oi, os_ = oracle_snapshot(nodes, pods, sid, 1, per_pod=per_pod)
        assert_same_idx(gi, oi, f"n={n} p={p} strategy={sid} per_pod={per_pod}")
        assert_same_f64(gs, os_)


def test_all_equal_nodes_first_wins():
    nodes = synth.all_equal_nodes(6000)
    nodes.ready[:5] = 0
    nodes.pod_count[5] = 110                                  # full: ineligible under capacity
    nodes.cpu_usage_pct[5] = nodes.mem_usage_pct[5] = 50.0
    _, pods = synth.config2(10, 1000)
    for sid in (0, 1):
        idx, _ = device_snapshot(nodes, pods, sid, 1)
        assert (idx == 6).all()
        idx, _ = device_snapshot(nodes, pods, sid, 0)
        assert (idx == 6).all()        # node 5 scores lower (pc=110), so 6 still wins
    idx, _ = device_snapshot(nodes, pods, 2, 0)
    assert (idx == 5).all()            # round_robin = MOST pods (scheduler.py:544-549)


def test_edge_cases():
    nodes, pods = synth.config2(50, 40)
    empty_pods = PodTable.from_columns([], [])
    idx, score = device_snapshot(nodes, empty_pods, 0, 1)
    assert idx.size == 0 and score.size == 0
    empty_nodes = NodeTable.from_columns(alloc_cpu_milli=[], alloc_mem_bytes=[], pod_count=[],
                                         max_pods=[], ready=[], cpu_usage_pct=[], mem_usage_pct=[])
    for mode in (0, 1):
        idx, score = device_snapshot(empty_nodes, pods, 0, mode)
        assert (idx == -1).all() and (score == -1.0).all()
    nodes.ready[:] = 0
    for mode in (0, 1):
        idx, score = device_snapshot(nodes, pods, 1, mode)
        assert (idx == -1).all() and (score == -1.0).all()
    # nothing fits: every pod larger than every node
    nodes, pods = synth.config2(300, 64)
    pods.req_cpu_milli[:] = 10 ** 9
    idx, score = device_snapshot(nodes, pods, 0, 1)
    assert (idx == -1).all() and (score == -1.0).all()
    # sentinel: score <= -1 never wins (scheduler.py:529,547); NaN usage never wins; max_pods == 0
    specs = [(341, 110, 155.0, 155.0, 1), (0, 0, 0.0, 0.0, 1), (5, 110, float("nan"), 1.0, 1),
             (223, 110, 101.36363636363636, 101.36363636363636, 1)]
    t = table_from_specs(specs)
    one = PodTable.from_columns([1], [1])
    for sid in (0, 1, 2):
        gi, gs = device_snapshot(t, one, sid, 0)
        oi, os_ = oracle_snapshot(t, one, sid, 0)
        assert_same_idx(gi, oi)
        assert_same_f64(gs, os_)
        gi, gs = device_snapshot(t, one, sid, 1)
        oi, os_ = oracle_snapshot(t, one, sid, 1)
        assert_same_idx(gi, oi)
        assert_same_f64(gs, os_)


def test_many_distinct_thresholds_take_the_wide_path():
    """More than 32000 distinct free-cpu / free-memory values: the packed 15-bit rank table
    cannot be used and place_kernel falls back to two 32-bit rank compares per pair.  Also a
    table with distinct values but below the limit, and continuous (non-A3) usage values."""
    rng = np.random.Generator(np.random.PCG64(31))
    for n, p in ((40_000, 3000), (6_000, 2000)):
        mp = np.full(n, 110, dtype=np.int64)
        pc = rng.integers(0, 110, n)
        nodes = NodeTable.from_columns(alloc_cpu_milli=rng.integers(4000, 96000, n),
                                       alloc_mem_bytes=rng.integers(16, 768, n) << 30, pod_count=pc,
                                       max_pods=mp, ready=(rng.random(n) > 0.02).astype(np.uint8),
                                       cpu_usage_pct=rng.random(n) * 90.0, mem_usage_pct=rng.random(n) * 90.0)
        pods = PodTable.from_columns(rng.integers(50, 60000, p), rng.integers(64, 400 * 1024, p) << 20,
                                     rng.integers(0, 3, p))
        for sid in (0, 1):
            gi, gs = device_snapshot(nodes, pods, sid, 1)
            oi, os_ = oracle_snapshot(nodes, pods, sid, 1)
            assert_same_idx(gi, oi, f"n={n} strategy={sid}")
            assert_same_f64(gs, os_)
        gi, gs = device_snapshot(nodes, pods, 0, 1, per_pod=True)
        oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
        assert_same_idx(gi, oi)
        assert_same_f64(gs, os_)


def test_large_values_and_negative_requests():
    rng = np.random.Generator(np.random.PCG64(77))
    n, p = 500, 2000
    mp = rng.integers(1, 10 ** 6, n)
    pc = (rng.random(n) * mp * 1.2).astype(np.int64)
    u = synth.usage_a3(pc, mp)
    nodes = NodeTable.from_columns(alloc_cpu_milli=rng.integers(1, 2 ** 50, n),
                                   alloc_mem_bytes=rng.integers(1, 2 ** 52, n), pod_count=pc,
                                   max_pods=mp, ready=np.ones(n, np.uint8), cpu_usage_pct=u,
                                   mem_usage_pct=u * rng.random(n))
    pods = PodTable.from_columns(rng.integers(-5, 2 ** 49, p), rng.integers(-5, 2 ** 51, p))
    for sid in (0, 1, 2):
        gi, gs = device_snapshot(nodes, pods, sid, 1)
        oi, os_ = oracle_snapshot(nodes, pods, sid, 1)
        assert_same_idx(gi, oi)
        assert_same_f64(gs, os_)


def test_full_size_properties_config4_shard():
    """50k nodes x 160k pods (a config-4 shard): too big for the oracle at full P, so check
    size-independent properties plus the oracle on a seeded subsample."""
    import torch
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = synth.config4(50_000, 160_000)
    gi, gs = device_snapshot(nodes, pods, 0, 1)
    gi2, gs2 = device_snapshot(nodes, pods, 0, 1)
    assert np.array_equal(gi, gi2) and np.array_equal(gs.view(np.uint64), gs2.view(np.uint64))
    # pod-permutation equivariance
    perm = np.random.Generator(np.random.PCG64(5)).permutation(len(pods))
    shuffled = PodTable.from_columns(pods.req_cpu_milli[perm], pods.req_mem_bytes[perm])
    pi, ps = device_snapshot(nodes, shuffled, 0, 1)
    assert np.array_equal(pi, gi[perm]) and np.array_equal(ps.view(np.uint64), gs[perm].view(np.uint64))
    # the chosen node fits, and reports its own score
    dn = placement.nodes_to_device(nodes, "cuda:0")
    score, fc, fm, el = (t.cpu().numpy() for t in placement.node_scores(dn, 0, 1))
    ok = gi >= 0
    assert ok.all()
    assert el[gi].all()
    assert (pods.req_cpu_milli / 1000.0 <= fc[gi]).all()
    assert (pods.req_mem_bytes / 1024 / 1024 / 1024 <= fm[gi]).all()
    assert np.array_equal(score[gi].view(np.uint64), gs.view(np.uint64))
    # oracle on a subsample
    sub = np.random.Generator(np.random.PCG64(6)).choice(len(pods), 1500, replace=False)
    sp = PodTable.from_columns(pods.req_cpu_milli[sub], pods.req_mem_bytes[sub])
    oi, os_ = oracle_snapshot(nodes, sp, 0, 1)
    assert_same_idx(gi[sub], oi)
    assert_same_f64(gs[sub], os_)


def test_session_host_api_matches_device_api():
    from k8s_llm_scheduler_b200.placement import Session
    nodes, pods = synth.config5(3000, 20_000, seed=42)
    uniform = PodTable.from_columns(pods.req_cpu_milli, pods.req_mem_bytes)   # no per-pod strategy ids
    with Session(0, max_nodes=4000, max_pods=32_768) as s:
        for mode in (0, 1):
            hi, hs = s.snapshot(nodes, uniform, 1, mode)
            di, ds = device_snapshot(nodes, uniform, 1, mode)
            assert_same_idx(hi, di)
            assert_same_f64(hs, ds)
        hi, hs = s.snapshot(nodes, pods, 0, 1)                      # per-pod strategy ids travel too
        oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
        assert_same_idx(hi, oi)
        assert_same_f64(hs, os_)
        h2d, d2h = s.last_copy_bytes()
        assert h2d == 6 * 8 * 3000 + 3000 + 3 * 8 * 20_000 and d2h == 16 * 20_000
        s.set_nodes(nodes)
        ri, rs = s.snapshot(None, uniform, 0, 1)                   # resident table reused
        oi, os_ = oracle_snapshot(nodes, uniform, 0, 1)
        assert_same_idx(ri, oi)
        from k8s_llm_scheduler_b200 import PlacementError
        with pytest.raises(PlacementError):
            s.snapshot(nodes, PodTable.from_columns(np.zeros(40_000), np.zeros(40_000)), 0, 1)


@pytest.mark.parametrize("n", [1000, 2500, 12_000])
def test_session_resident_table_reuses_its_node_plan(n):
    """kp_session_set_nodes keeps the table on the device; snapshots with nodes=None that follow with the same
    strategy / filter reuse the node plan of the snapshot before (no plan kernels).  Every answer must stay
    the oracle's, whatever happens in between: other pods, another strategy, a greedy call on the same
    workspace, a new table."""
    from k8s_llm_scheduler_b200 import placement
    from k8s_llm_scheduler_b200.placement import Session
    nodes, pods = synth.config5(n, 30_000, seed=50 + n)
    other, _ = synth.config5(n, 1, seed=51 + n)
    a = PodTable.from_columns(pods.req_cpu_milli[:20_000], pods.req_mem_bytes[:20_000])
    b = PodTable.from_columns(pods.req_cpu_milli[20_000:], pods.req_mem_bytes[20_000:])
    lib = placement._lib.load()
    with Session(0, max_nodes=n, max_pods=30_000) as s:
        s.set_nodes(nodes)

        def check(table, part, strat, mode=1, per_pod=False):
            before = placement.launch_count()
            gi, gs = s.snapshot(None, part, strat, mode)
            oi, os_ = oracle_snapshot(table, part, strat, mode, per_pod=per_pod)
            assert_same_idx(gi, oi)
            assert_same_f64(gs, os_)
            return placement.launch_count() - before

        first = check(nodes, a, 0)
        again = check(nodes, a, 0)
        assert again < first and again == 1                    # only the placement kernel ran
        assert check(nodes, b, 0) == 1                         # another pod count: the plan's place depends on the nodes alone
        assert check(nodes, b, 1) == first                     # another strategy: its table is not in the plan
        assert check(nodes, b, 1) == 1
        assert check(nodes, a, 1, mode=0) >= 1                 # the reference's own filter has no plan
        assert check(nodes, a, 1) == first
        with_ids = PodTable.from_columns(a.req_cpu_milli, a.req_mem_bytes, pods.strategy_ids[:20_000])
        check(nodes, with_ids, 0, per_pod=True)                # per-pod strategies: never reused ...
        assert check(nodes, a, 1) == first                     # ... and they invalidate what was there
        assert check(nodes, a, 1) == 1
        s.greedy(None, PodTable.from_columns(a.req_cpu_milli[:3000], a.req_mem_bytes[:3000]), 1, 1)   # takes the workspace
        assert check(nodes, a, 1) == first
        s.set_nodes(other)                                     # a new table
        assert check(other, a, 1) == first
        assert check(other, b, 1) == 1
        gi, _ = s.snapshot(nodes, a, 1, 1)                     # table passed with the call: planned afresh
        assert_same_idx(gi, oracle_snapshot(nodes, a, 1, 1)[0])
    assert lib is not None


@pytest.mark.parametrize("n,pinned", [(3000, False), (9000, True)])
def test_session_uploads_many_pods_in_chunks_behind_the_scoring(n, pinned):
    """More than a million pods with one strategy: the session sends the pod columns in chunks on a second stream
    and scores each chunk as it lands, all against the plan of the first chunk.  Same answers as one call."""
    import torch
    from k8s_llm_scheduler_b200.placement import Session
    p = 1_150_001
    nodes, pods = synth.config5(n, p, seed=60 + n)
    uniform = PodTable.from_columns(pods.req_cpu_milli, pods.req_mem_bytes)
    want = device_snapshot(nodes, uniform, 0, 1)
    head = PodTable.from_columns(pods.req_cpu_milli[:40_000], pods.req_mem_bytes[:40_000])
    oi, os_ = oracle_snapshot(nodes, head, 0, 1)
    assert_same_idx(want[0][:40_000], oi)
    with Session(0, max_nodes=n, max_pods=p) as s:
        if pinned:
            hp = (torch.as_tensor(pods.req_cpu_milli).pin_memory(), torch.as_tensor(pods.req_mem_bytes).pin_memory(), None)
            out = (torch.empty(p, dtype=torch.int64).pin_memory(), torch.empty(p, dtype=torch.float64).pin_memory())
            for _ in range(2):
                out[0].fill_(-7)
                s.snapshot(nodes, hp, 0, 1, out=out)
                assert_same_idx(out[0].numpy(), want[0])
                assert_same_f64(out[1].numpy(), want[1])
        else:
            gi, gs = s.snapshot(nodes, uniform, 0, 1)
            assert_same_idx(gi, want[0])
            assert_same_f64(gs, want[1])
        assert s.last_copy_bytes() == (6 * 8 * n + n + 16 * p, 16 * p)


@pytest.mark.parametrize("n", [700, 5000, 11_000])
def test_place_again_reuses_the_plan_in_the_workspace(n):
    """kp_snapshot_place_again: other pods against the same node table, only the placement kernel runs."""
    import torch
    from k8s_llm_scheduler_b200 import placement, PlacementError
    nodes, pods = synth.config5(n, 24_000, seed=70 + n)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    ws = torch.empty(placement.snapshot_workspace_bytes(n, 24_000, False), dtype=torch.uint8, device="cuda")
    placement.snapshot_place(dn, rc[:9000].contiguous(), rm[:9000].contiguous(), 1, 1, workspace=ws)
    before = placement.launch_count()
    gi, gs = placement.snapshot_place_again(dn, rc[9000:].contiguous(), rm[9000:].contiguous(), 1, ws)
    torch.cuda.synchronize()
    assert placement.launch_count() - before == 1
    rest = PodTable.from_columns(pods.req_cpu_milli[9000:], pods.req_mem_bytes[9000:])
    oi, os_ = oracle_snapshot(nodes, rest, 1, 1)
    assert_same_idx(gi.cpu().numpy(), oi)
    assert_same_f64(gs.cpu().numpy(), os_)
    with pytest.raises(PlacementError):                            # per-pod strategy plans are not reusable
        lib = placement._lib.load()
        nst, device, _ = placement._node_struct(dn)
        pst, p, strat, _ = placement._pod_struct(rc, rm, torch.as_tensor(pods.strategy_ids).cuda(), device)
        placement.check(lib.kp_snapshot_place_again(ctypes.byref(nst), ctypes.byref(pst), 0, 1, gi.data_ptr(), gs.data_ptr(),
                                                    ws.data_ptr(), ws.numel(), None), "kp_snapshot_place_again")


def test_session_inputs_that_need_conversion_stay_alive_for_the_call():
    """Columns of the 'wrong' dtype or layout (bool ready, int32 counts, Python lists, strided views, float
    tensors) are converted to temporaries; the temporaries must live until kp_session_snapshot has returned."""
    import gc
    import torch
    from k8s_llm_scheduler_b200.placement import Session
    nodes, pods = synth.config5(4000, 50_000, seed=43)
    uniform = PodTable.from_columns(pods.req_cpu_milli, pods.req_mem_bytes)
    want = oracle_snapshot(nodes, uniform, 0, 1)
    c = nodes.columns()
    odd = {"alloc_cpu_milli": c["alloc_cpu_milli"].tolist(), "alloc_mem_bytes": np.repeat(c["alloc_mem_bytes"], 2)[::2],
           "pod_count": c["pod_count"].astype(np.int32), "max_pods": torch.as_tensor(c["max_pods"]).to(torch.int32),
           "ready": c["ready"].astype(bool), "cpu_usage_pct": c["cpu_usage_pct"].astype(np.longdouble),
           "mem_usage_pct": torch.as_tensor(c["mem_usage_pct"])}
    odd_pods = (pods.req_cpu_milli.astype(np.float64), torch.as_tensor(pods.req_mem_bytes).to(torch.float64), None)
    with Session(0, max_nodes=4096, max_pods=65_536) as s:
        for _ in range(3):
            junk = [np.full(50_000, -1, np.int64) for _ in range(8)]      # churn the allocator between calls
            gi, gs = s.snapshot(odd, odd_pods, 0, 1)
            del junk
            gc.collect()
            assert_same_idx(gi, want[0])
            assert_same_f64(gs, want[1])


@pytest.mark.parametrize("n,p", [(1000, 30_000), (8192, 5_000), (9000, 5_000), (3, 1), (700, 257), (500, 400_000)])
def test_session_pinned_buffers_are_used_in_place(n, p):
    """Pinned host tables: small tables with one strategy are read / written by the kernels
    directly (no staging copies); everything else takes the copy path.  Same answers either way."""
    import torch
    from k8s_llm_scheduler_b200.placement import Session
    nodes, pods = synth.config5(n, p, seed=77)
    uniform = PodTable.from_columns(pods.req_cpu_milli, pods.req_mem_bytes)
    hn = {k: torch.as_tensor(getattr(nodes, k)).pin_memory() for k in nodes.columns()}
    hp = (torch.as_tensor(pods.req_cpu_milli).pin_memory(), torch.as_tensor(pods.req_mem_bytes).pin_memory(), None)
    hp_sid = hp[:2] + (torch.as_tensor(pods.strategy_ids).pin_memory(),)
    out = (torch.full((p,), -7, dtype=torch.int64).pin_memory(), torch.full((p,), -7.0, dtype=torch.float64).pin_memory())
    with Session(0, max_nodes=n, max_pods=p) as s:
        for strat in (0, 1, 2):
            for mode in (0, 1):
                out[0].fill_(-7)
                s.snapshot(hn, hp, strat, mode, out=out)
                oi, os_ = oracle_snapshot(nodes, uniform, strat, mode)
                assert_same_idx(out[0].numpy(), oi)
                assert_same_f64(out[1].numpy(), os_)
                assert s.last_copy_bytes() == (6 * 8 * n + n + 16 * p, 16 * p)
        s.snapshot(hn, hp_sid, 0, 1, out=out)                       # per-pod ids: copy path
        oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
        assert_same_idx(out[0].numpy(), oi)
        assert_same_f64(out[1].numpy(), os_)
        # pinned inputs, pageable outputs and the other way round
        gi, gs = s.snapshot(hn, hp, 0, 1)
        oi, os_ = oracle_snapshot(nodes, uniform, 0, 1)
        assert_same_idx(gi, oi)
        s.snapshot(nodes, uniform, 0, 1, out=out)
        assert_same_idx(out[0].numpy(), oi)
        assert_same_f64(out[1].numpy(), os_)


@pytest.mark.parametrize("n,p,per_pod", [(1000, 20_000, False), (12_000, 30_000, False), (9_000, 20_000, True)])
def test_snapshot_captures_into_a_cuda_graph(n, p, per_pod):
    """kp_snapshot_place is stream-ordered with caller-owned memory (no allocation, no synchronisation,
    side streams joined back): it can be captured once and replayed on new table contents."""
    import torch
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = synth.config5(n, p, seed=51)
    dev = torch.device("cuda:0")
    dn = placement.nodes_to_device(nodes, dev)
    rc, rm = torch.as_tensor(pods.req_cpu_milli).to(dev), torch.as_tensor(pods.req_mem_bytes).to(dev)
    strat = torch.as_tensor(pods.strategy_ids).to(dev) if per_pod else 0
    ws = torch.empty(placement.snapshot_workspace_bytes(n, p, per_pod), dtype=torch.uint8, device=dev)
    out = (torch.empty(p, dtype=torch.int64, device=dev), torch.empty(p, dtype=torch.float64, device=dev))

    def step():
        placement.snapshot_place(dn, rc, rm, strat, 1, workspace=ws, out=out)

    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        step()                                           # first call outside the capture (one-time attribute set-up)
        torch.cuda.synchronize()
        with torch.cuda.graph(graph, stream=side):
            step()
    # new requests and a changed node table in the same buffers, then replay
    _, pods2 = synth.config5(n, p, seed=52)
    rc.copy_(torch.as_tensor(pods2.req_cpu_milli))
    rm.copy_(torch.as_tensor(pods2.req_mem_bytes))
    nodes.pod_count[::3] = np.minimum(nodes.pod_count[::3] + 1, np.maximum(nodes.max_pods[::3], 0))
    dn["pod_count"].copy_(torch.as_tensor(nodes.pod_count))
    out[0].fill_(-9)
    graph.replay()
    torch.cuda.synchronize()
    pods2.strategy_ids = pods.strategy_ids
    oi, os_ = oracle_snapshot(nodes, pods2, 0, 1, per_pod=per_pod)
    assert_same_idx(out[0].cpu().numpy(), oi)
    assert_same_f64(out[1].cpu().numpy(), os_)


def test_workspace_errors_are_reported():
    import torch
    from k8s_llm_scheduler_b200 import PlacementError, placement
    nodes, pods = synth.config2(100, 100)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    with pytest.raises(ValueError):
        placement.snapshot_place(dn, rc, rm, 0, 1, workspace=torch.empty(16, dtype=torch.uint8, device="cuda"))
    lib = placement._lib.load()
    nst, _, _ = placement._node_struct(dn)
    pst, *_ = placement._pod_struct(rc, rm, 0, rc.device)
    out_i, out_s = torch.empty(100, dtype=torch.int64, device="cuda"), torch.empty(100, dtype=torch.float64, device="cuda")
    ws = torch.empty(1024, dtype=torch.uint8, device="cuda")
    st = lib.kp_snapshot_place(ctypes.byref(nst), ctypes.byref(pst), 0, 1, out_i.data_ptr(), out_s.data_ptr(),
                               ws.data_ptr(), 1024, None)
    assert st == 2 and b"workspace" in lib.kp_last_error()


def test_full_size_config5_snapshot():
    """BASELINE config 5 at full size (20k heterogeneous nodes x 2M pods, per-pod strategies, ~30 %
    infeasible): properties at full size, exact parity with the oracle on a seeded subsample."""
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = synth.config5(20_000, 2_000_000)
    gi, gs = device_snapshot(nodes, pods, 0, 1, per_pod=True)
    infeasible = (gi < 0).mean()
    assert 0.2 < infeasible < 0.4
    assert (gs[gi < 0] == -1.0).all()
    dn = placement.nodes_to_device(nodes, "cuda:0")
    for sid in (0, 1, 2):
        score, fc, fm, el = (t.cpu().numpy() for t in placement.node_scores(dn, sid, 1))
        mine = (pods.strategy_ids == sid) & (gi >= 0)
        g = gi[mine]
        assert el[g].all()
        assert (pods.req_cpu_milli[mine] / 1000.0 <= fc[g]).all()
        assert (pods.req_mem_bytes[mine] / 1024 / 1024 / 1024 <= fm[g]).all()
        assert np.array_equal(score[g].view(np.uint64), gs[mine].view(np.uint64))
    sub = np.random.Generator(np.random.PCG64(8)).choice(len(pods), 3000, replace=False)
    sp = PodTable.from_columns(pods.req_cpu_milli[sub], pods.req_mem_bytes[sub], pods.strategy_ids[sub])
    oi, os_ = oracle_snapshot(nodes, sp, 0, 1, per_pod=True)
    assert_same_idx(gi[sub], oi)
    assert_same_f64(gs[sub], os_)


def test_full_size_config4_whole():
    """BASELINE config 4 whole on one GPU: 50k nodes x 10M pods, resource_balanced."""
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = synth.config4(50_000, 10_000_000)
    gi, gs = device_snapshot(nodes, pods, 0, 1)
    assert (gi >= 0).all()
    dn = placement.nodes_to_device(nodes, "cuda:0")
    score, fc, fm, el = (t.cpu().numpy() for t in placement.node_scores(dn, 0, 1))
    assert el[gi].all()
    assert (pods.req_cpu_milli / 1000.0 <= fc[gi]).all()
    assert (pods.req_mem_bytes / 1024 / 1024 / 1024 <= fm[gi]).all()
    assert np.array_equal(score[gi].view(np.uint64), gs.view(np.uint64))
    sub = np.random.Generator(np.random.PCG64(9)).choice(len(pods), 1000, replace=False)
    sp = PodTable.from_columns(pods.req_cpu_milli[sub], pods.req_mem_bytes[sub])
    oi, os_ = oracle_snapshot(nodes, sp, 0, 1)
    assert_same_idx(gi[sub], oi)
    assert_same_f64(gs[sub], os_)
