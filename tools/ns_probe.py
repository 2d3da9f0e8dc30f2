"""Where a node-sharded snapshot step goes, kernel group by kernel group: W ranks emulated on one GPU (lock step),
rank 0's calls timed with CUDA events.  python tools/ns_probe.py [world] [nodes] [pods]   (GPU box)"""
import ctypes
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from k8s_llm_scheduler_b200 import _lib, dist as kd, placement, synth  # noqa: E402

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20_000
p = int(sys.argv[3]) if len(sys.argv) > 3 else 2_000_000
nodes, pods = synth.config5(n, p)
dev = torch.device("cuda:0")
lib = _lib.load()
cuts = [kd.shard_bounds(n, r, world) for r in range(world)]
group = kd.NodeShardedPlacer.local_group([dev] * world, max_pods=p, max_local_nodes=max(b - a for a, b in cuts))
import os  # noqa: E402
if os.environ.get("NS_PRUNE"):                      # optional first-feasible probe in every rank's scan
    for g in group:
        g.set_prune_depth(int(os.environ["NS_PRUNE"]))
    print(f"(first-feasible probe, depth {os.environ['NS_PRUNE']})")
full = placement.nodes_to_device(nodes, dev)
slices = [kd.slice_nodes(full, a, b) for a, b in cuts]
rc, rm = torch.as_tensor(pods.req_cpu_milli).to(dev), torch.as_tensor(pods.req_mem_bytes).to(dev)
sid = torch.as_tensor(pods.strategy_ids).to(dev)
for per_pod in (True, False):
    strat = sid if per_pod else 0
    acc = {}
    for it in range(6):
        ev = {}

        def timed(name, fn):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            ev[name] = (a, b)

        for g, (lo, _) in zip(group, cuts):
            f = lambda g=g, lo=lo: g.keys(slices[g.rank], lo, rc, rm, strat, 1)
            timed("plan+publish", f) if g.rank == 0 else f()
        lib.kp_profile_enable(1)
        for g in group:
            f = lambda g=g: g.scan(slices[g.rank], rc, rm, strat, 1)
            if g.rank == 0:
                timed("gather+gpos+groups+place+signal", f)
                fa, fb = ctypes.c_float(), ctypes.c_float()
                lib.kp_profile_last_ms(ctypes.byref(fa), ctypes.byref(fb))
                plan_ms, place_ms = fa.value, fb.value
            else:
                f()
        lib.kp_profile_enable(0)
        for g in group:
            f = lambda g=g: g.reduce()
            timed("reduce (local peers)", f) if g.rank == 0 else f()
        torch.cuda.synchronize()
        if it >= 2:
            for k, (a, b) in ev.items():
                acc.setdefault(k, []).append(a.elapsed_time(b) * 1e3)
            acc.setdefault("  of which plan (+groups)", []).append(plan_ms * 1e3)
            acc.setdefault("  of which place kernel", []).append(place_ms * 1e3)
    print(f"world {world} emulated, {n} nodes ({cuts[0][1]} per rank) x {p} pods, {'per-pod strategies' if per_pod else 'one strategy'}:")
    for k, v in acc.items():
        print(f"   {k:34s} {np.median(v):9.1f} us")
