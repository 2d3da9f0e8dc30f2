"""One rank's share of a node-sharded config 5: a 2 500-node slice x 2 M pods, per-pod strategies -- the two
placement kernels (nodes in registers vs pods in registers) side by side.  python tools/slice_probe.py [n_slice]"""
import ctypes
import os
import statistics
import subprocess
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
if len(sys.argv) > 2:
    import torch
    from k8s_llm_scheduler_b200 import _lib, placement, synth
    from k8s_llm_scheduler_b200.dist import slice_nodes
    n_slice = int(sys.argv[1])
    nodes, pods = synth.config5(20_000, 2_000_000)
    dev = torch.device("cuda:0")
    dn = slice_nodes(placement.nodes_to_device(nodes, dev), 0, n_slice)
    rc, rm = torch.as_tensor(pods.req_cpu_milli).to(dev), torch.as_tensor(pods.req_mem_bytes).to(dev)
    sid = torch.as_tensor(pods.strategy_ids).to(dev)
    lib = _lib.load()
    for label, strat in (("per-pod", sid), ("uniform", 0)):
        lib.kp_profile_enable(1)
        plan, place = [], []
        for i in range(8):
            placement.snapshot_place(dn, rc, rm, strat, 1)
            a, b = ctypes.c_float(), ctypes.c_float()
            lib.kp_profile_last_ms(ctypes.byref(a), ctypes.byref(b))
            if i >= 3:
                plan.append(a.value); place.append(b.value)
        t = statistics.median(place)
        print(f"{sys.argv[2]:10s} {label:8s} slice {n_slice}: plan {statistics.median(plan)*1e3:7.1f} us  place {t*1e3:8.1f} us  "
              f"{n_slice * 2e6 / t / 1e9:6.2f} Tpairs/s", flush=True)
else:
    n_slice = sys.argv[1] if len(sys.argv) > 1 else "2500"
    for tag, env in (("registers", {}), ("shared-mem", {"KP_SMALL_OLD": "1"})):
        subprocess.run([sys.executable, __file__, n_slice, tag], env={**os.environ, **env}, check=True)
