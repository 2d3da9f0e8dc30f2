"""kp_session_snapshot on config 2 with pinned host buffers: table sent per call vs resident table (plan reused),
for the in-place and the staged-copy forms (KP_SESSION_DIRECT = 3 / 2 / 0).  GPU box: python tools/session_probe.py"""
import os
import statistics
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402
import bench  # noqa: E402
from k8s_llm_scheduler_b200 import placement  # noqa: E402
from k8s_llm_scheduler_b200.tables import NODE_COLUMNS  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
nodes, pods, meta = bench.make_workload(name)
p = len(pods)
hn = {k: torch.as_tensor(getattr(nodes, k)).pin_memory() for k in NODE_COLUMNS}
hp = (torch.as_tensor(pods.req_cpu_milli).pin_memory(), torch.as_tensor(pods.req_mem_bytes).pin_memory(), None)
ho = (torch.empty(p, dtype=torch.int64).pin_memory(), torch.empty(p, dtype=torch.float64).pin_memory())
for direct in ("3", "2", "0"):
    os.environ["KP_SESSION_DIRECT"] = direct
    with placement.Session(0, max_nodes=len(nodes), max_pods=p) as sess:
        for order in (("resident", None), ("per-call", hn), ("resident", None), ("per-call", hn)):
            label, table = order
            sess.set_nodes(hn)
            for _ in range(10):
                sess.snapshot(table, hp, meta["strategy"], meta["filter"], out=ho)
            ts = []
            for _ in range(200):
                t = time.perf_counter()
                sess.snapshot(table, hp, meta["strategy"], meta["filter"], out=ho)
                ts.append(time.perf_counter() - t)
            print(f"direct={direct} {label:9s} median {statistics.median(ts)*1e6:7.1f} us  min {min(ts)*1e6:7.1f} us", flush=True)
