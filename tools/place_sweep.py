#!/usr/bin/env python3
"""Experiment driver: time plan + place kernels (kp_profile events) for the BASELINE shapes.
Usage: KP_LIB_PATH=<variant.so> python tools/place_sweep.py [c2 c4shard c5 ...]"""
import ctypes, os, statistics, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from k8s_llm_scheduler_b200 import _lib, placement
import bench

def run(name, steps=5):
    nodes, pods, meta = bench.make_workload(name)
    dev = torch.device("cuda", 0)
    dn = placement.nodes_to_device(nodes, dev)
    rc = torch.as_tensor(pods.req_cpu_milli).to(dev); rm = torch.as_tensor(pods.req_mem_bytes).to(dev)
    strat = torch.as_tensor(pods.strategy_ids).to(dev) if meta["per_pod"] else meta["strategy"]
    lib = _lib.load(); lib.kp_profile_enable(1)
    ws = torch.empty(placement.snapshot_workspace_bytes(len(nodes), len(pods), meta["per_pod"]), dtype=torch.uint8, device=dev)
    plan, place = [], []
    for i in range(steps + 2):
        placement.snapshot_place(dn, rc, rm, strat, meta["filter"], workspace=ws)
        a, b = ctypes.c_float(), ctypes.c_float()
        lib.kp_profile_last_ms(ctypes.byref(a), ctypes.byref(b))
        if i >= 2: plan.append(a.value); place.append(b.value)
    pm = statistics.median(place)
    print(f"{os.environ.get('KP_LIB_PATH','default'):40s} {name:8s} plan {statistics.median(plan)*1e3:8.1f} us  place {pm*1e3:9.1f} us  "
          f"{len(nodes)*len(pods)/pm/1e9:8.2f} Tpairs/s in-kernel", flush=True)

for w in (sys.argv[1:] or ["c2", "c4shard", "c5"]):
    run(w)


def loop_avg(name, iters=300):
    """Average device time per call over many back-to-back calls (event resolution is ~2 us)."""
    nodes, pods, meta = bench.make_workload(name)
    dev = torch.device("cuda", 0)
    dn = placement.nodes_to_device(nodes, dev)
    rc = torch.as_tensor(pods.req_cpu_milli).to(dev); rm = torch.as_tensor(pods.req_mem_bytes).to(dev)
    strat = torch.as_tensor(pods.strategy_ids).to(dev) if meta["per_pod"] else meta["strategy"]
    _lib.load().kp_profile_enable(0)
    ws = torch.empty(placement.snapshot_workspace_bytes(len(nodes), len(pods), meta["per_pod"]), dtype=torch.uint8, device=dev)
    out = (torch.empty(len(pods), dtype=torch.int64, device=dev), torch.empty(len(pods), dtype=torch.float64, device=dev))
    for _ in range(10):
        placement.snapshot_place(dn, rc, rm, strat, meta["filter"], workspace=ws, out=out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        placement.snapshot_place(dn, rc, rm, strat, meta["filter"], workspace=ws, out=out)
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / iters
    print(f"{name}: {us:.2f} us per call over {iters} back-to-back calls = {len(nodes)*len(pods)/us/1e6:.2f} Tpairs/s", flush=True)


if os.environ.get("KP_SWEEP_LOOP"):
    loop_avg("c2")
