"""kp_session_decide: ONE decision for ONE pod (the reference's own call shape, scheduler.py:598 -> :521-559) against
the CPU oracle -- node index exact, score bit for bit.  Tables of up to 2048 rows are scored where they lie (one
kernel reading pinned host memory, answer polled); larger ones go through the batched path."""
import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200._lib import FILTER_CAPACITY, FILTER_READY
from k8s_llm_scheduler_b200.tables import NodeTable, PodTable
from tests.helpers import bits, oracle_snapshot

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def session():
    from k8s_llm_scheduler_b200.placement import Session
    s = Session(max_nodes=8192, max_pods=64)
    yield s
    s.close()


@pytest.mark.parametrize("n", [1, 2, 3, 31, 32, 33, 63, 64, 65, 700, 1024, 1025, 2047, 2048, 2049, 5000])
def test_decide_matches_oracle(session, n):
    nodes, pods = synth.config5(n, 6, seed=4000 + n)
    for filt in (FILTER_READY, FILTER_CAPACITY):
        for sid in (0, 1, 2, 7):
            oi, os_ = oracle_snapshot(nodes, pods, sid, filt)
            for p in range(len(pods)):
                i, sc = session.decide(nodes, sid, filt, int(pods.req_cpu_milli[p]), int(pods.req_mem_bytes[p]))
                assert i == oi[p], (n, filt, sid, p)
                assert bits(np.float64(sc)) == bits(os_[p]), (n, filt, sid, p)


def test_decide_ties_sentinels_and_empty_tables(session):
    n = 100
    same = NodeTable.from_columns(alloc_cpu_milli=[8000] * n, alloc_mem_bytes=[16 << 30] * n, pod_count=[5] * n,
                                  max_pods=[110] * n, ready=[0] * 7 + [1] * (n - 7), cpu_usage_pct=[10.0] * n,
                                  mem_usage_pct=[20.0] * n)
    for sid in (0, 1, 2):
        assert session.decide(same, sid, FILTER_READY)[0] == 7             # first Ready node in list order wins the tie
        assert session.decide(same, sid, FILTER_CAPACITY, 1000, 1 << 30)[0] == 7
        assert session.decide(same, sid, FILTER_CAPACITY, 64_000, 1 << 30) == (-1, -1.0)      # fits nowhere
    none = NodeTable.from_columns(alloc_cpu_milli=[8000] * 3, alloc_mem_bytes=[16 << 30] * 3, pod_count=[1, 2, 3],
                                  max_pods=[110] * 3, ready=[0, 0, 0], cpu_usage_pct=[1.0] * 3, mem_usage_pct=[1.0] * 3)
    assert session.decide(none, 0, FILTER_READY) == (-1, -1.0)
    # scores that cannot beat the reference's -1 sentinel with a strict '>' (scheduler.py:529,547)
    low = NodeTable.from_columns(alloc_cpu_milli=[8000] * 2, alloc_mem_bytes=[16 << 30] * 2, pod_count=[1, 1],
                                 max_pods=[110] * 2, ready=[1, 1], cpu_usage_pct=[250.0, float("nan")],
                                 mem_usage_pct=[250.0, 1.0])
    assert session.decide(low, 1, FILTER_READY) == (-1, -1.0)               # -300 and NaN: neither is > -1
    empty = NodeTable.from_columns(alloc_cpu_milli=[], alloc_mem_bytes=[], pod_count=[], max_pods=[], ready=[],
                                   cpu_usage_pct=[], mem_usage_pct=[])
    assert session.decide(empty, 0, FILTER_READY) == (-1, -1.0)


def test_decide_call_after_call_never_returns_a_stale_answer(session):
    """The answer is polled from one pinned slot: every call must see ITS kernel's answer, whatever ran before
    (other tables, other sizes, batched snapshots on the same session in between)."""
    rng = np.random.default_rng(5)
    tables = []
    for k in range(12):
        n = int(rng.integers(1, 300))
        nodes, pods = synth.config5(n, 1, seed=6000 + k)
        oi, os_ = oracle_snapshot(nodes, pods, k % 3, FILTER_CAPACITY)
        tables.append((nodes, pods, k % 3, int(oi[0]), os_[0]))
    big_nodes, big_pods = synth.config5(3000, 50, seed=9)
    for it in range(600):
        nodes, pods, sid, want_i, want_s = tables[int(rng.integers(0, len(tables)))]
        i, sc = session.decide(nodes, sid, FILTER_CAPACITY, int(pods.req_cpu_milli[0]), int(pods.req_mem_bytes[0]))
        assert i == want_i and bits(np.float64(sc)) == bits(want_s), it
        if it % 97 == 0:
            session.snapshot(big_nodes, big_pods, 0, FILTER_CAPACITY)


def test_decide_rejects_what_the_session_cannot_hold(session):
    from k8s_llm_scheduler_b200._lib import PlacementError
    nodes, _ = synth.config5(9000, 1, seed=1)
    with pytest.raises(PlacementError):
        session.decide(nodes, 0, FILTER_READY)
    with pytest.raises(PlacementError):
        session.decide(synth.config5(10, 1, seed=1)[0], 0, 5)
