#!/usr/bin/env python3
"""Where does the end-to-end time of kp_session_snapshot go?  (GPU box only)
Times the host API on the c2 shape (or --nodes/--pods) with pinned host tables:
  * staging copies vs pinned buffers used in place (KP_SESSION_DIRECT=0/1),
  * through the Python Session wrapper vs the bare C call with prebuilt structs,
  * node table passed every call vs kept resident.
Prints one line per variant (median / min µs over --iters calls)."""
import argparse
import ctypes
import os
import statistics
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from k8s_llm_scheduler_b200 import placement, synth  # noqa: E402
from k8s_llm_scheduler_b200.tables import NODE_COLUMNS  # noqa: E402


def timeit(fn, iters, warm=20):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        t = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t) * 1e6)
    return statistics.median(ts), min(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1000)
    ap.add_argument("--pods", type=int, default=100_000)
    ap.add_argument("--strategy", type=int, default=1)
    ap.add_argument("--iters", type=int, default=300)
    ap.add_argument("--direct", default="0,1,2,3", help="KP_SESSION_DIRECT masks to time")
    ap.add_argument("--windows", default="", help="flow-control windows (CTAs in flight) to time, e.g. 0,32,96@5")
    ap.add_argument("--cfgs", default="", help="place configurations to force, e.g. 3,4,5")
    a = ap.parse_args()
    nodes, pods = synth.config2(a.nodes, a.pods)
    hn = {k: torch.as_tensor(getattr(nodes, k)).pin_memory() for k in NODE_COLUMNS}
    hp = (torch.as_tensor(pods.req_cpu_milli).pin_memory(), torch.as_tensor(pods.req_mem_bytes).pin_memory(), None)
    ho = (torch.empty(a.pods, dtype=torch.int64).pin_memory(), torch.empty(a.pods, dtype=torch.float64).pin_memory())
    # what the host link alone costs for this much data (copy engine, one direction at a time)
    dbuf = torch.empty(2 * a.pods, dtype=torch.int64, device="cuda")
    hbuf = torch.empty(2 * a.pods, dtype=torch.int64).pin_memory()

    def h2d():
        dbuf.copy_(hbuf, non_blocking=True)
        torch.cuda.synchronize()

    def d2h():
        hbuf.copy_(dbuf, non_blocking=True)
        torch.cuda.synchronize()

    for name, fn in (("H2D", h2d), ("D2H", d2h)):
        med, mn = timeit(fn, a.iters)
        print(f"link only: one {name} copy of {16 * a.pods} B + sync : median {med:8.1f} us  min {mn:8.1f} us")
    ref = None
    ap_cfgs = [c for c in a.cfgs.split(",") if c]
    for item in ap_cfgs:                       # "cfg" or "cfg:resident"
        cfg, _, res = item.partition(":")
        os.environ["KP_SESSION_DIRECT"] = "3"
        os.environ["KP_FORCE_PLACE_CFG"] = cfg
        if res:
            os.environ["KP_PLACE_RESIDENT"] = res
        sess = placement.Session(0, max_nodes=a.nodes, max_pods=a.pods)
        med, mn = timeit(lambda: sess.snapshot(hn, hp, a.strategy, 1, out=ho), a.iters)
        print(f"direct=3 cfg={cfg} resident={res or '-'} python wrapper, nodes every call : median {med:8.1f} us  min {mn:8.1f} us")
        sess.close()
        os.environ.pop("KP_PLACE_RESIDENT", None)
    os.environ.pop("KP_FORCE_PLACE_CFG", None)
    for item in [w for w in a.windows.split(",") if w]:     # "window" or "window@cfg"
        win, _, cfg = item.partition("@")
        os.environ["KP_SESSION_DIRECT"] = "3"
        os.environ["KP_FLOW_WINDOW"] = win
        if cfg:
            os.environ["KP_FORCE_PLACE_CFG"] = cfg
        sess = placement.Session(0, max_nodes=a.nodes, max_pods=a.pods)
        med, mn = timeit(lambda: sess.snapshot(hn, hp, a.strategy, 1, out=ho), a.iters)
        print(f"direct=3 flow window={win} cfg={cfg or '-'} python wrapper, nodes every call : median {med:8.1f} us  min {mn:8.1f} us")
        if ref is None:
            os.environ["KP_SESSION_DIRECT"] = "0"
            s0 = placement.Session(0, max_nodes=a.nodes, max_pods=a.pods)
            ref = s0.snapshot(hn, hp, a.strategy, 1)[0].copy()
            s0.close()
        assert np.array_equal(ref, ho[0].numpy()), "paced variant disagrees"
        sess.close()
        os.environ.pop("KP_FORCE_PLACE_CFG", None)
    os.environ.pop("KP_FLOW_WINDOW", None)
    for direct in a.direct.split(","):
        os.environ["KP_SESSION_DIRECT"] = direct
        sess = placement.Session(0, max_nodes=a.nodes, max_pods=a.pods)
        lib = sess._lib
        med, mn = timeit(lambda: sess.snapshot(hn, hp, a.strategy, 1, out=ho), a.iters)
        print(f"direct={direct} python wrapper, nodes every call : median {med:8.1f} us  min {mn:8.1f} us")
        idx = ho[0].numpy().copy()
        if ref is None:
            ref = idx
        assert np.array_equal(ref, idx), "variants disagree"
        nst, keep1 = sess._node_struct(hn)
        pst, keep2 = sess._pod_struct(hp)
        ip, sp = ho[0].data_ptr(), ho[1].data_ptr()

        def bare():
            st = lib.kp_session_snapshot(sess._h, ctypes.byref(nst), ctypes.byref(pst), a.strategy, 1, ip, sp)
            assert st == 0

        med, mn = timeit(bare, a.iters)
        print(f"direct={direct} bare C call,    nodes every call : median {med:8.1f} us  min {mn:8.1f} us")
        sess.set_nodes(hn)

        def resident():
            st = lib.kp_session_snapshot(sess._h, None, ctypes.byref(pst), a.strategy, 1, ip, sp)
            assert st == 0

        med, mn = timeit(resident, a.iters)
        print(f"direct={direct} bare C call,    resident nodes   : median {med:8.1f} us  min {mn:8.1f} us")
        sess.close()
    print("ok")


if __name__ == "__main__":
    main()
