"""Small end-to-end case touching every kernel family, for compute-sanitizer runs."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from k8s_llm_scheduler_b200 import placement, synth
from oracle import c_oracle

def snap(nodes, pods, strat, mode, per_pod):
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc = torch.as_tensor(pods.req_cpu_milli).cuda(); rm = torch.as_tensor(pods.req_mem_bytes).cuda()
    s = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else strat
    i, sc = placement.snapshot_place(dn, rc, rm, s, mode)
    oi, os_ = c_oracle.snapshot(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes, strategy=strat,
                                strategy_ids=pods.strategy_ids if per_pod else None, filter_mode=mode, threads=4)
    assert np.array_equal(i.cpu().numpy(), oi) and np.array_equal(sc.cpu().numpy().view(np.uint64), os_.view(np.uint64))

def greedy(nodes, pods, strat, mode, per_pod):
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc = torch.as_tensor(pods.req_cpu_milli).cuda(); rm = torch.as_tensor(pods.req_mem_bytes).cuda()
    s = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else strat
    g = placement.greedy_place(dn, rc, rm, s, mode)
    o = c_oracle.greedy(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes, strategy=strat,
                        strategy_ids=pods.strategy_ids if per_pod else None, filter_mode=mode)
    assert np.array_equal(g[0].cpu().numpy(), o[0]) and np.array_equal(g[2].cpu().numpy(), o[2])

n1, p1 = synth.config5(700, 3000, seed=1)        # small-table plan (cluster count kernel), NSPLIT place
snap(n1, p1, 1, 1, False); snap(n1, p1, 0, 1, True); snap(n1, p1, 0, 0, True)
n2, p2 = synth.config5(9000, 2500, seed=2)       # large-table plan (sorts), node segments + merge
snap(n2, p2, 0, 1, False); snap(n2, p2, 0, 1, True)
n3, p3 = synth.config5(1500, 2600, seed=3)       # greedy fast path (256 threads) + per-pod strategies
greedy(n3, p3, 0, 1, False); greedy(n3, p3, 0, 1, True)
greedy(n3, p3, 0, 0, False)                      # fp64 kernel (ready filter)
import os
os.environ["KP_GREEDY_CLUSTER"] = "2"
greedy(n3, p3, 0, 1, True)                       # cluster + DSMEM exchange
os.environ.pop("KP_GREEDY_CLUSTER", None)
# ingest kernels (strings -> columns), incl. rows the device defers to the host
from k8s_llm_scheduler_b200 import ingest, tables
reqs = [[{"cpu": "250m", "memory": "256Mi"}], [{"cpu": "0.1", "memory": "1Gi"}, {"cpu": "200m", "memory": "12x"}], [], [{"cpu": "1e3m"}]]
out = ingest.ingest_pod_requests(reqs[:1] + reqs[2:3] * 300, exact=False)
assert out.table.req_cpu_milli[0] == 250
rows = [dict(name=f"n{i}", cpu="7910m", memory="32780300Ki", pods="110", pod_count=i % 111) for i in range(500)]
assert np.array_equal(ingest.ingest_nodes(rows).table.alloc_cpu_milli, tables.NodeTable.from_allocatable(rows).alloc_cpu_milli)
# host session with pinned buffers (in-place reads / writes, paced)
hn = {k: torch.as_tensor(getattr(n1, k)).pin_memory() for k in tables.NODE_COLUMNS}
hp = (torch.as_tensor(p1.req_cpu_milli).pin_memory(), torch.as_tensor(p1.req_mem_bytes).pin_memory(), None)
ho = (torch.empty(len(p1), dtype=torch.int64).pin_memory(), torch.empty(len(p1), dtype=torch.float64).pin_memory())
with placement.Session(0, max_nodes=len(n1), max_pods=len(p1)) as sess:
    sess.snapshot(hn, hp, 1, 1, out=ho)
    one = sess.decide(n1, 1, 1, int(p1.req_cpu_milli[0]), int(p1.req_mem_bytes[0]))     # one decision, host-memory kernel
oi, _ = c_oracle.snapshot(n1.columns(), p1.req_cpu_milli, p1.req_mem_bytes, strategy=1, filter_mode=1, threads=4)
assert np.array_equal(ho[0].numpy(), oi)
assert one[0] == oi[0]
torch.cuda.synchronize()
print("sanitize case ok")
