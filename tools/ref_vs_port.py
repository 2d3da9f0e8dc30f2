#!/usr/bin/env python3
"""BUILD CONTAINER ONLY (needs /root/reference): is bench.py's CPU baseline -- the CPython port oracle/py_oracle.py --
representative of the reference's own code?  Times, on ONE core of this host and on the same synthetic tables,
  (a) the unmodified reference `HuggingFaceClient._fallback_decision` (imported as tests/golden/make_golden.py does),
  (b) the port with the reference's own filter (Ready only), record for record the same work,
  (c) the port with the capacity filter (what bench.py times, because that is what the GPU path computes).
Prints one line per node count; the numbers are quoted in BASELINE.md.  Test infrastructure: touches oracle/."""
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests" / "golden"))
import make_golden  # noqa: E402
from k8s_llm_scheduler_b200 import synth  # noqa: E402
from oracle import py_oracle as po  # noqa: E402

sched = make_golden.import_reference()
for n, strategy, name in ((1000, 1, "least_loaded"), (50_000, 0, "resource_balanced")):
    nodes, pods = (synth.config2(n, 2000) if n == 1000 else synth.config4(n, 60))
    recs = po.nodes_from_arrays(**nodes.columns())
    ref_nodes = [sched.NodeMetrics(name=r.name, cpu_usage_percent=r.cpu_usage_percent, memory_usage_percent=r.memory_usage_percent,
                                   available_cpu=r.available_cpu, available_memory=r.available_memory, pod_count=r.pod_count,
                                   max_pods=r.max_pods, labels={}, taints=[], conditions=r.conditions) for r in recs]
    specs = po.pods_from_arrays(pods.req_cpu_milli, pods.req_mem_bytes)
    sched.CONFIG.setdefault("fallback", {})["strategy"] = name
    p = len(specs)

    def best(fn, reps=3):
        ts = []
        for _ in range(reps):
            t = time.perf_counter(); fn(); ts.append(time.perf_counter() - t)
        return min(ts)

    t_ref = best(lambda: [sched.HuggingFaceClient._fallback_decision(None, ref_nodes, "x") for _ in range(p)])
    t_ready = best(lambda: po.snapshot_place(recs, specs, name, po.FILTER_READY))
    t_cap = best(lambda: po.snapshot_place(recs, specs, name, po.FILTER_CAPACITY))
    same = [sched.HuggingFaceClient._fallback_decision(None, ref_nodes, "x").selected_node for _ in range(3)] == \
           [recs[i].name if i >= 0 else "" for i in po.snapshot_place(recs, specs[:3], name, po.FILTER_READY)[0]]
    print(f"{n} nodes x {p} decisions, {name}, one core: reference {p * n / t_ref / 1e6:.2f} M pairs/s | port, Ready filter "
          f"{p * n / t_ready / 1e6:.2f} M pairs/s | port, capacity filter {p * n / t_cap / 1e6:.2f} M pairs/s | same node: {same}")
