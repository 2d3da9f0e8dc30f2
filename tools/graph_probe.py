"""Does the snapshot path capture into a CUDA graph, and what does a replayed step cost?  (GPU box)
python tools/graph_probe.py [c2|c4shard|c5]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from k8s_llm_scheduler_b200 import placement  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
nodes, pods, meta = bench.make_workload(name)
dev = torch.device("cuda:0")
dn = placement.nodes_to_device(nodes, dev)
rc = torch.as_tensor(pods.req_cpu_milli).to(dev)
rm = torch.as_tensor(pods.req_mem_bytes).to(dev)
strat = torch.as_tensor(pods.strategy_ids).to(dev) if meta["per_pod"] else meta["strategy"]
ws = torch.empty(placement.snapshot_workspace_bytes(len(nodes), len(pods), meta["per_pod"]), dtype=torch.uint8, device=dev)
out = (torch.empty(len(pods), dtype=torch.int64, device=dev), torch.empty(len(pods), dtype=torch.float64, device=dev))


def step():
    placement.snapshot_place(dn, rc, rm, strat, meta["filter"], workspace=ws, out=out)


for _ in range(3):
    step()
torch.cuda.synchronize()
want = out[0].clone()
side = torch.cuda.Stream()
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(side):
    step()
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=side):
        step()
out[0].fill_(-5)
g.replay()
torch.cuda.synchronize()
print("graph replay reproduces the direct call:", bool(torch.equal(out[0], want)))


def avg_us(fn, iters=300):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def single_us(fn, iters=50):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(iters):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts))


print(f"{name}: back-to-back  direct {avg_us(step):8.2f} us   graph {avg_us(g.replay):8.2f} us")
print(f"{name}: flushed, each step timed alone  direct {single_us(step):8.2f} us   graph {single_us(g.replay):8.2f} us")
