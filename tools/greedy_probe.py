import sys, time
sys.path.insert(0, '/root/repo')
import torch
from k8s_llm_scheduler_b200 import placement, synth
for (n, p, per_pod) in ((5000, 200000, False), (20000, 50000, False), (20000, 50000, True), (3000, 200000, True)):
    nodes, pods = synth.config5(n, p, seed=5) if per_pod or n == 20000 else synth.config3(n, p)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc = torch.as_tensor(pods.req_cpu_milli).cuda(); rm = torch.as_tensor(pods.req_mem_bytes).cuda()
    strat = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else 0
    placement.greedy_place(dn, rc[:1000].contiguous(), rm[:1000].contiguous(), strat[:1000].contiguous() if per_pod else 0, 1)
    torch.cuda.synchronize(); t = time.perf_counter()
    placement.greedy_place(dn, rc, rm, strat, 1)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"greedy n={n} p={p} per_pod={per_pod}: {dt*1e3:.1f} ms = {dt/p*1e6:.3f} us/pod = {p/dt/1e6:.2f} M pods/s", flush=True)
