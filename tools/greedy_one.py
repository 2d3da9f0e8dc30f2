"""One greedy placement for profiling: python tools/greedy_one.py [c3|c5|c5u] [n_pods] (GPU box)."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402
from k8s_llm_scheduler_b200 import placement, synth  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "c3"
p = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
if kind == "c5":
    nodes, pods = synth.config5(20_000, p)
    strat = torch.as_tensor(pods.strategy_ids).cuda()
elif kind == "c5u":                       # config-5 nodes and pods, one strategy for all pods
    nodes, pods = synth.config5(20_000, p)
    strat = 0
else:
    nodes, pods = synth.config3(5000, p)
    strat = 0
dn = placement.nodes_to_device(nodes, "cuda:0")
rc = torch.as_tensor(pods.req_cpu_milli).cuda()
rm = torch.as_tensor(pods.req_mem_bytes).cuda()
for _ in range(2):
    torch.cuda.synchronize()
    t = time.perf_counter()
    placement.greedy_place(dn, rc, rm, strat, 1)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
print(f"ok {kind} {p} pods: {dt * 1e3:.1f} ms, {p / dt / 1e6:.3f} M pods/s")
