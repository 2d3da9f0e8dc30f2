"""Small end-to-end case touching every kernel family, for compute-sanitizer runs."""
import sys
sys.path.insert(0, '/root/repo')
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
torch.cuda.synchronize()
print("sanitize case ok")
