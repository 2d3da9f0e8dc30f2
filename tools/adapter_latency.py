"""Latency of ONE fallback decision through the drop-in adapter vs the CPython loop it replaces
(oracle's literal restatement of scheduler.py:521-559), for several cluster sizes.  (GPU box)"""
import statistics
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from k8s_llm_scheduler_b200.adapter import FallbackPlacer  # noqa: E402
from k8s_llm_scheduler_b200.tables import NodeTable  # noqa: E402
from oracle import py_oracle  # noqa: E402


def med_us(fn, iters):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(iters):
        t = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t) * 1e6)
    return statistics.median(ts)


placer = FallbackPlacer({"fallback": {"strategy": "resource_balanced"}}, max_nodes=65536, max_pods=16)
for n in (3, 100, 1000, 10000, 50000):
    nodes = [py_oracle.node_row(f"n{i}", "32", "128Gi", "110", i % 111) for i in range(n)]
    iters = 200 if n <= 1000 else 20
    gpu = med_us(lambda: placer.fallback_decision(nodes, "x"), iters)
    cpu = med_us(lambda: py_oracle.fallback_decision(nodes, "x", "resource_balanced"), iters)
    from k8s_llm_scheduler_b200._lib import FILTER_READY  # noqa: E402
    pack = med_us(lambda: placer._node_columns(nodes, ready_only=True), iters)
    st = placer._node_columns(nodes, ready_only=True)
    bare = med_us(lambda: placer.session.decide(st, 0, FILTER_READY), iters)
    a, b = placer.fallback_decision(nodes, "x"), py_oracle.fallback_decision(nodes, "x", "resource_balanced")
    assert a.selected_node == b.selected_node
    print(f"{n:6d} nodes: adapter {gpu:9.1f} us (of which list -> columns {pack:9.1f} us, kp_session_decide {bare:7.1f} us)   "
          f"CPython loop {cpu:9.1f} us")
placer.close()
