"""The first-feasible probe (kp_snapshot_place_pruned) beside the full table scan on the BASELINE shapes and on a
cluster whose first fit lies anywhere: step time, pods left for the scan, evaluated vs effective pairs.  (GPU box)"""
import statistics
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import bench  # noqa: E402
from k8s_llm_scheduler_b200 import placement  # noqa: E402
from k8s_llm_scheduler_b200.tables import NodeTable, PodTable  # noqa: E402


def scattered(n, p, seed=1, tight=0.5):
    rng = np.random.default_rng(seed)
    nodes = NodeTable.from_columns(
        alloc_cpu_milli=rng.integers(500, 64_000, n), alloc_mem_bytes=rng.integers(1 << 28, 256 << 30, n),
        pod_count=rng.integers(0, 110, n), max_pods=rng.choice([30, 110, 110, 250], n),
        ready=(rng.random(n) > 0.03).astype(np.uint8), cpu_usage_pct=np.round(rng.random(n) * 100, 2),
        mem_usage_pct=np.round(rng.random(n) * 100, 2))
    pods = PodTable.from_columns(rng.integers(1, int(64_000 * tight), p), rng.integers(1, int((256 << 30) * tight), p))
    return nodes, pods


def timed(fn, reps=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return statistics.median(ms)


dev = torch.device("cuda:0")
cases = [(name,) + bench.make_workload(name)[:2] + (bench.make_workload(name)[2],) for name in ("c2", "c5", "c4")]
cases.append(("scattered 20k x 2M", *scattered(20_000, 2_000_000), {"per_pod": False, "strategy": 0}))
cases.append(("scattered-tight 20k x 2M", *scattered(20_000, 2_000_000, tight=0.95), {"per_pod": False, "strategy": 0}))
for name, nodes, pods, meta in cases:
    n, p = len(nodes), len(pods)
    dn = placement.nodes_to_device(nodes, dev)
    rc, rm = torch.as_tensor(pods.req_cpu_milli).to(dev), torch.as_tensor(pods.req_mem_bytes).to(dev)
    strat = torch.as_tensor(pods.strategy_ids).to(dev) if meta["per_pod"] else meta["strategy"]
    ws = torch.empty(placement.snapshot_workspace_bytes(n, p, True), dtype=torch.uint8, device=dev)
    out = (torch.empty(p, dtype=torch.int64, device=dev), torch.empty(p, dtype=torch.float64, device=dev))
    full = timed(lambda: placement.snapshot_place(dn, rc, rm, strat, 1, workspace=ws, out=out))
    ref = out[0].clone()
    print(f"{name:26s} {n:6d} nodes x {p:9d} pods   full scan {full:9.3f} ms ({n * p / full / 1e9:7.2f} T pairs/s)", flush=True)
    for depth in (1, 8, 32, 64):
        t = timed(lambda: placement.snapshot_place_pruned(dn, rc, rm, strat, 1, depth, workspace=ws, out=out))
        _, _, left = placement.snapshot_place_pruned(dn, rc, rm, strat, 1, depth, workspace=ws, out=out, scanned=True)
        same = bool(torch.equal(out[0], ref))
        print(f"{'':26s} probe depth {depth:3d}: {t:9.3f} ms  x{full / t:7.1f}   pods left for the scan {sum(left):9d} "
              f"({100.0 * sum(left) / p:5.2f} %)   effective {n * p / t / 1e9:9.2f} T pairs/s   same answers: {same}", flush=True)
