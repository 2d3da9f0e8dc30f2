#!/usr/bin/env python3
"""BASELINE config 5 in its node-sharded form: every rank scores ALL pods against its contiguous
slice of the node table, then the per-pod winner under (score desc, index asc) is two NCCL
all-reduces (MAX over scores, MIN over the indices that attain it).  Strong scaling: the job
(20 000 nodes x 2 000 000 pods, mixed strategies) is fixed, the node slice shrinks with N.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      tools/node_sharded_bench.py [--steps K] [--warmup W] [--pods P]

Device time per step (CUDA events on the launching stream, max over ranks); rank 0 prints one JSON
line and checks a seeded subsample against the CPU oracle (tools/ may use oracle/ as a checker)."""
import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from k8s_llm_scheduler_b200 import dist as kdist, placement, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--nodes", type=int, default=20_000)
    ap.add_argument("--pods", type=int, default=2_000_000)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nodes, pods = synth.config5(a.nodes, a.pods)                 # same seed on every rank: same job
    dn = placement.nodes_to_device(nodes, dev)
    rc = torch.as_tensor(pods.req_cpu_milli).to(dev)
    rm = torch.as_tensor(pods.req_mem_bytes).to(dev)
    sid = torch.as_tensor(pods.strategy_ids).to(dev)
    lo, hi = kdist.shard_bounds(len(nodes), rank, world)
    mine = kdist.slice_nodes(dn, lo, hi)

    def step():
        return kdist.snapshot_node_sharded(mine, lo, rc, rm, sid, 1)

    for _ in range(max(a.warmup, 3)):
        out = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    evs = []
    for _ in range(a.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = step()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    tot = torch.tensor([sum(x.elapsed_time(y) for x, y in evs)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    ms = float(tot.item()) / a.steps
    if rank == 0:
        from oracle import c_oracle
        sub = np.random.Generator(np.random.PCG64(3)).choice(a.pods, 2000, replace=False)
        oi, os_ = c_oracle.snapshot(nodes.columns(), pods.req_cpu_milli[sub], pods.req_mem_bytes[sub],
                                    strategy_ids=pods.strategy_ids[sub], filter_mode=1, threads=16)
        gi, gs = out[0].cpu().numpy()[sub], out[1].cpu().numpy()[sub]
        ok = bool(np.array_equal(gi, oi) and np.array_equal(gs.view(np.uint64), os_.view(np.uint64)))
        print(json.dumps({"metric": "pod_x_node_scores_per_sec", "value": a.nodes * a.pods / (ms * 1e-3),
                          "unit": "pairs/s", "n_gpus": world, "steps": a.steps, "ms_per_step": ms,
                          "scaling": "strong", "pods_per_sec": a.pods / (ms * 1e-3),
                          "config": {"workload": f"c5 node-sharded: {a.nodes} nodes / {world} ranks x {a.pods} pods, "
                                                 "mixed strategies, capacity filter, 2 NCCL all-reduces per step"},
                          "parity_vs_oracle_sample": ok}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
