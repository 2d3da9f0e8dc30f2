"""One snapshot placement of a BASELINE shape, for profiling: python tools/snap_one.py [c2|c4|c4shard|c5] [probe depth]
(GPU box).  With a probe depth the call is kp_snapshot_place_pruned."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402
import bench  # noqa: E402
from k8s_llm_scheduler_b200 import placement  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c5"
nodes, pods, meta = bench.make_workload(name)
dn = placement.nodes_to_device(nodes, "cuda:0")
rc = torch.as_tensor(pods.req_cpu_milli).cuda()
rm = torch.as_tensor(pods.req_mem_bytes).cuda()
strat = torch.as_tensor(pods.strategy_ids).cuda() if meta["per_pod"] else meta["strategy"]
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for _ in range(3):
    if depth > 0:
        placement.snapshot_place_pruned(dn, rc, rm, strat, meta["filter"], depth)
    else:
        placement.snapshot_place(dn, rc, rm, strat, meta["filter"])
torch.cuda.synchronize()
print("ok", name)
