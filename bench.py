#!/usr/bin/env python3
"""Benchmark of the fallback placement path (BASELINE.json metric: pod x node scores/sec
and pods placed/sec), own arm and reference arm.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c4shard|c5|c5greedy]
  python bench.py --impl reference ...        # the reference's CPU algorithm on host cores
  torchrun --nproc-per-node N bench.py --gpus N ...   # one rank per GPU, NCCL

A "step" is one pass of the hot path over one batch of synthetic input:
  * default workload c2 = BASELINE configs[1]: least_loaded snapshot scoring, 1 000 nodes x
    100 000 pending pods per GPU, capacity filter.  With N > 1 GPUs every rank scores its
    own 100 000-pod shard against the node table broadcast from rank 0 (weak scaling, no
    data-path collective).
  * `value` times the device path with the tables already resident in HBM (CUDA events on
    the launching stream, L2 flushed between steps, max over ranks).
  * `e2e` times the reference-facing host API (kp_session_snapshot): pinned host tables in,
    H2D copies, kernels, D2H of the results, all inside the timed region.
  * `roofline` is for the dominant kernel (place_kernel), timed live with CUDA events.
  * `cpu_baseline` / `--impl reference` time the CPU restatement of the reference's loop
    (oracle/py_oracle.py, same CPython data structures) on the box's host cores over a
    bounded sample of the same workload.  Only these two legs touch oracle/.

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "pod_x_node_scores_per_sec"
UNIT = "pairs/s"
STRATEGY_NAMES = {0: "resource_balanced", 1: "least_loaded", 2: "round_robin"}


# --------------------------------------------------------------------------------------
# workloads; pods_per_gpu fixed as N grows (weak scaling)
# --------------------------------------------------------------------------------------
def make_workload(name: str, rank: int):
    from k8s_llm_scheduler_b200 import synth
    seed_shift = 1000 * rank              # every rank scores a different pod shard
    if name == "c2":
        nodes, _ = synth.config2(1000, 1)
        _, pods = synth.config2(1000, 100_000, seed=2 + seed_shift)
        return nodes, pods, dict(mode="snapshot", strategy=1, filter=1, per_pod=False,
                                 desc="c2: least_loaded snapshot, 1000 nodes x 100000 pods per GPU, capacity filter")
    if name == "c4shard":
        nodes, _ = synth.config4(50_000, 1)
        _, pods = synth.config4(50_000, 1_250_000, seed=4 + seed_shift)
        return nodes, pods, dict(mode="snapshot", strategy=0, filter=1, per_pod=False,
                                 desc="c4 shard: resource_balanced snapshot, 50000 nodes x 1250000 pods per GPU "
                                      "(10M pods / 8), capacity filter")
    if name == "c4":
        nodes, _ = synth.config4(50_000, 1)
        _, pods = synth.config4(50_000, 10_000_000, seed=4 + seed_shift)
        return nodes, pods, dict(mode="snapshot", strategy=0, filter=1, per_pod=False,
                                 desc="c4: resource_balanced snapshot, 50000 nodes x 10000000 pods per GPU, capacity filter")
    if name == "c5":
        nodes, _ = synth.config5(20_000, 1)
        _, pods = synth.config5(20_000, 2_000_000, seed=5 + seed_shift)
        return nodes, pods, dict(mode="snapshot", strategy=0, filter=1, per_pod=True,
                                 desc="c5 (snapshot form): mixed strategies, 20000 heterogeneous nodes x 2000000 pods "
                                      "per GPU, 30% infeasible, capacity filter")
    if name == "c3":
        nodes, pods = synth.config3(5000, 1_000_000, seed=3 + seed_shift)
        return nodes, pods, dict(mode="greedy", strategy=0, filter=1, per_pod=False,
                                 desc="c3: resource_balanced greedy commit, 5000 nodes x 1000000 pods per GPU, "
                                      "capacity filter (replicas at N > 1)")
    if name == "c5greedy":
        nodes, _ = synth.config5(20_000, 1)
        _, pods = synth.config5(20_000, 2_000_000, seed=5 + seed_shift)
        return nodes, pods, dict(mode="greedy", strategy=0, filter=1, per_pod=True,
                                 desc="c5 (greedy form): mixed strategies, 20000 heterogeneous nodes x 2000000 pods "
                                      "per GPU, 30% infeasible, sequential commit (replicas at N > 1)")
    raise SystemExit(f"unknown workload {name}")


# --------------------------------------------------------------------------------------
# clocks during the timed region
# --------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, pw, reasons = [], [], [], set()
        for ts, line in self.rows:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------
# CPU port of the reference loop, on all host cores (snapshot) or one core (greedy)
# --------------------------------------------------------------------------------------
_W = {}


def _cpu_worker(span):
    from oracle import py_oracle as po
    lo, hi = span
    nodes, pods, names, mode = _W["nodes"], _W["pods"], _W["names"], _W["filter"]
    idx, _ = po.snapshot_place(nodes, pods[lo:hi], names if isinstance(names, str) else names[lo:hi], mode)
    return idx


def host_cores() -> int:
    return len(os.sched_getaffinity(0))


def cpu_port_run(nodes, pods, meta, budget_s: float, calibrate=True):
    """Time oracle/py_oracle.py on a bounded sample.  Returns dict(value, cores, sample, idx, n_sample)."""
    import multiprocessing as mp
    from oracle import py_oracle as po
    n = len(nodes)
    recs = po.nodes_from_arrays(**nodes.columns())
    names = ([po.strategy_name(s) for s in pods.strategy_ids] if meta["per_pod"]
             else po.strategy_name(meta["strategy"]))
    cores = host_cores() if meta["mode"] == "snapshot" else 1
    # calibrate: a few pods on one core
    probe = po.pods_from_arrays(pods.req_cpu_milli[:4], pods.req_mem_bytes[:4])
    t = time.perf_counter()
    po.snapshot_place(recs, probe, names if isinstance(names, str) else names[:4], meta["filter"])
    per_pod = max((time.perf_counter() - t) / 4, 1e-7)
    s = int(max(cores, min(len(pods), budget_s / per_pod * cores)))
    sample = po.pods_from_arrays(pods.req_cpu_milli[:s], pods.req_mem_bytes[:s])
    if meta["mode"] == "greedy":
        work = po.nodes_from_arrays(**nodes.columns())
        t = time.perf_counter()
        idx, _ = po.greedy_place(work, sample, names if isinstance(names, str) else names[:s], meta["filter"])
        el = time.perf_counter() - t
    else:
        _W.update(nodes=recs, pods=sample, names=names, filter=meta["filter"])
        cuts = np.linspace(0, s, cores + 1).astype(int)
        spans = [(int(cuts[i]), int(cuts[i + 1])) for i in range(cores)]
        ctx = mp.get_context("fork")
        with ctx.Pool(cores) as pool:
            pool.map(_cpu_worker, [(0, 1)] * cores)          # spin the workers up untimed
            t = time.perf_counter()
            parts = pool.map(_cpu_worker, spans, chunksize=1)
            el = time.perf_counter() - t
        idx = [i for part in parts for i in part]
    return dict(value=s * n / el, pods_per_s=s / el, seconds=el, cores=cores, n_sample=s, idx=np.array(idx),
                sample=f"first {s} pods x all {n} nodes of the workload, CPython port of the reference loop "
                       f"(oracle/py_oracle.py), {cores} process(es), {el:.1f} s")


# --------------------------------------------------------------------------------------
# reference arm
# --------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nodes, pods, meta = make_workload(args.workload, 0)
    per_step = max(1.0, min(12.0, 150.0 / (args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_port_run(nodes, pods, meta, per_step)
    vals, secs, last = [], 0.0, None
    for _ in range(args.steps):
        last = cpu_port_run(nodes, pods, meta, per_step)
        vals.append(last["value"])
        secs += last["seconds"]
    value = float(np.mean(vals))
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "pods_per_sec": value / len(nodes),
        "config": {"workload": meta["desc"], "nodes": len(nodes), "strategy": STRATEGY_NAMES[meta["strategy"]],
                   "sample_pods_per_step": last["n_sample"]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["cores"], "kind": "port",
                         "sample": last["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference is pure Python with a K8s/HF import chain that cannot travel to the GPU box; "
                "this arm times its CPU restatement (pinned to reference-generated vectors) on the host cores",
    }
    print(json.dumps(out))


# --------------------------------------------------------------------------------------
# own arm
# --------------------------------------------------------------------------------------
def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            d = json.loads(p.read_text())
            return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, 1965.0, "fallback"


def run_ours(args):
    import torch
    import torch.distributed as dist
    from k8s_llm_scheduler_b200 import _lib, placement
    from k8s_llm_scheduler_b200.build import build_library
    from k8s_llm_scheduler_b200.tables import NODE_COLUMNS, NodeTable

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    build_library()
    lib = _lib.load()

    def barrier():
        if world > 1:
            dist.barrier()

    nodes, pods, meta = make_workload(args.workload, rank)
    n, p = len(nodes), len(pods)
    # node table: rank 0's copy is broadcast over NCCL/NVLink (replicated table, sharded pods)
    dn = placement.nodes_to_device(nodes, dev)
    if world > 1:
        for k in NODE_COLUMNS:
            dist.broadcast(dn[k], src=0)
    rc = torch.as_tensor(pods.req_cpu_milli).to(dev)
    rm = torch.as_tensor(pods.req_mem_bytes).to(dev)
    strat = torch.as_tensor(pods.strategy_ids).to(dev) if meta["per_pod"] else meta["strategy"]
    snapshot = meta["mode"] == "snapshot"
    if snapshot:
        ws = torch.empty(placement.snapshot_workspace_bytes(n, p, meta["per_pod"]), dtype=torch.uint8, device=dev)
        out = (torch.empty(p, dtype=torch.int64, device=dev), torch.empty(p, dtype=torch.float64, device=dev))

        def step():
            placement.snapshot_place(dn, rc, rm, strat, meta["filter"], workspace=ws, out=out)
    else:
        ws = torch.empty(placement.greedy_workspace_bytes(n, p), dtype=torch.uint8, device=dev)
        res = {}

        def step():
            res["out"] = placement.greedy_place(dn, rc, rm, strat, meta["filter"], workspace=ws)

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def flush():
        if not args.no_flush:
            flush_buf.zero_()

    hbm_peak, sm_max_mhz, peak_src = measured_peaks()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    import ctypes
    place_ms, plan_ms = [], []

    # timed region (device time, per-step CUDA events on the launching stream); no
    # instrumentation inside the steps, so the kernels chain with dependent launches
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    launches0 = placement.launch_count()
    evs = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    launches = placement.launch_count() - launches0
    # second pass of the same K steps with CUDA events around the plan and the place kernel
    # (kp_profile_*): the live duration of the dominant kernel for the roofline
    if snapshot and meta["filter"] == 1:
        lib.kp_profile_enable(1)
        for _ in range(args.steps):
            flush()
            step()
            fa, fb = ctypes.c_float(), ctypes.c_float()
            if lib.kp_profile_last_ms(ctypes.byref(fa), ctypes.byref(fb)) == 0:   # waits for this step
                plan_ms.append(fa.value)
                place_ms.append(fb.value)
        lib.kp_profile_enable(0)
        torch.cuda.synchronize()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    tot = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_ms = float(tot.item())
    ms_per_step = total_ms / args.steps
    pairs_per_step = float(world) * p * n
    value = pairs_per_step / (ms_per_step * 1e-3)
    clocks = sampler.stop(t0, t1) if rank == 0 else None

    # end-to-end through the host API: pinned host tables in, results out, every step
    sess = placement.Session(local, max_nodes=n, max_pods=p)
    hn = {k: torch.as_tensor(getattr(nodes, k)).pin_memory() for k in NODE_COLUMNS}
    hp = (torch.as_tensor(pods.req_cpu_milli).pin_memory(), torch.as_tensor(pods.req_mem_bytes).pin_memory(),
          torch.as_tensor(pods.strategy_ids).pin_memory() if meta["per_pod"] else None)
    ho = (torch.empty(p, dtype=torch.int64).pin_memory(), torch.empty(p, dtype=torch.float64).pin_memory())

    def e2e_step():
        if snapshot:
            sess.snapshot(hn, hp, meta["strategy"], meta["filter"], out=ho)
        else:
            sess.greedy(hn, hp, meta["strategy"], meta["filter"])

    e2e_steps = args.steps if snapshot else max(1, min(args.steps, 3))
    for _ in range(min(args.warmup, 3) if not snapshot else args.warmup):
        e2e_step()
    barrier()
    t = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    e2e_s = torch.tensor([(time.perf_counter() - t) / e2e_steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    h2d, d2h = sess.last_copy_bytes()
    e2e_value = pairs_per_step / float(e2e_s.item())
    e2e_idx = ho[0].numpy().copy() if snapshot else None
    sess.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # roofline of the dominant kernel
    roofline = None
    if place_ms:
        t_place = statistics.mean(place_ms) * 1e-3
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        clk = (clocks or {}).get("sm_mhz") or sm_max_mhz
        peak_issue = sm_count * 128 * sm_max_mhz * 1e6           # lane-ops/s: 4 schedulers x 32 lanes per SM-clock
        ach = 3.0 * p * n / t_place                              # 2 compares + 1 select per pair
        roofline = {
            "kernel": "place_kernel", "bound": "issue", "achieved": ach / 1e12, "peak": peak_issue / 1e12,
            "unit": "Tlane-op/s", "frac": ach / peak_issue, "traffic": None,
            "peak_source": f"{sm_count} SMs x 128 lanes x {sm_max_mhz:.0f} MHz (MEASURED_PEAKS.json sm_max_mhz, {peak_src})",
            "frac_at_sampled_clock": ach / (sm_count * 128 * clk * 1e6),
            "achieved_survey_4op": 4.0 * p * n / t_place / 1e12,
            "kernel_ms": t_place * 1e3, "plan_ms": statistics.mean(plan_ms),
            "kernel_share_of_step": statistics.mean(place_ms) / statistics.mean(step_ms),
            "kernel_timing": "CUDA events around place_kernel in a second pass of the same steps (instrumented)",
            "hbm": {"bound": "hbm", "achieved": (32.0 * p + 56.0 * n) / t_place / 1e9, "peak": hbm_peak,
                    "unit": "GB/s", "frac": (32.0 * p + 56.0 * n) / t_place / 1e9 / hbm_peak,
                    "note": "algorithmic HBM bytes are negligible for this kernel; issue-bound (SURVEY.md §8d regime X)"},
        }
        tr = ROOT / "profiles" / "traffic.json"
        if tr.exists():
            try:
                roofline["traffic"] = json.loads(tr.read_text()).get(args.workload)
            except Exception:  # noqa: BLE001
                pass

    cpu_baseline = None
    parity = None
    if world == 1 and not args.no_cpu:
        cb = cpu_port_run(nodes, pods, meta, args.cpu_seconds)
        cpu_baseline = {"value": cb["value"], "unit": UNIT, "cores": cb["cores"], "kind": "port",
                        "sample": cb["sample"], "pods_per_sec": cb["pods_per_s"]}
        if snapshot:
            got = e2e_idx[:cb["n_sample"]]
        else:
            got = res["out"][0][:cb["n_sample"]].cpu().numpy()
        parity = bool(np.array_equal(got, cb["idx"]))

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 scores, i32 rank compares", "data": "synthetic",
        "pods_per_sec": float(world) * p / (ms_per_step * 1e-3),
        "config": {"workload": meta["desc"], "nodes": n, "pods_per_gpu": p,
                   "strategy": "per-pod mix" if meta["per_pod"] else STRATEGY_NAMES[meta["strategy"]],
                   "filter": "capacity" if meta["filter"] == 1 else "ready", "mode": meta["mode"],
                   "l2": "NOT flushed (diagnostic run)" if args.no_flush else "flushed between timed steps (256 MiB write)",
                   "sharding": "pods sharded per rank, node table broadcast from rank 0 (NCCL)" if world > 1 else "single GPU"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": float(e2e_s.item()) * 1e3, "api": "kp_session_snapshot" if snapshot else "kp_session_greedy"},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
        "cpu_baseline": cpu_baseline, "parity_vs_cpu_sample": parity,
    }
    if world == 1 and args.extra:
        out["extra"] = run_extras(args, torch, placement, lib, dev)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def run_extras(args, torch, placement, lib, dev):
    """Secondary single-GPU numbers for the other BASELINE configs (not the headline)."""
    import ctypes
    extras = {}
    for name, steps in (("c4shard", 3), ("c4", 2), ("c5", 2), ("c3", 1), ("c5greedy", 1)):
        try:
            nodes, pods, meta = make_workload(name, 0)
            n, p = len(nodes), len(pods)
            dn = placement.nodes_to_device(nodes, dev)
            rc = torch.as_tensor(pods.req_cpu_milli).to(dev)
            rm = torch.as_tensor(pods.req_mem_bytes).to(dev)
            strat = torch.as_tensor(pods.strategy_ids).to(dev) if meta["per_pod"] else meta["strategy"]
            if meta["mode"] == "snapshot":
                lib.kp_profile_enable(1)
                fn = lambda: placement.snapshot_place(dn, rc, rm, strat, meta["filter"])  # noqa: E731
            else:
                fn = lambda: placement.greedy_place(dn, rc, rm, strat, meta["filter"])  # noqa: E731
            fn()
            torch.cuda.synchronize()
            ms, place = [], []
            for _ in range(steps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); fn(); b.record()
                torch.cuda.synchronize()
                ms.append(a.elapsed_time(b))
                if meta["mode"] == "snapshot":
                    fa, fb = ctypes.c_float(), ctypes.c_float()
                    if lib.kp_profile_last_ms(ctypes.byref(fa), ctypes.byref(fb)) == 0:
                        place.append(fb.value)
            lib.kp_profile_enable(0)
            t = statistics.mean(ms) * 1e-3
            e = {"workload": meta["desc"], "ms_per_step": t * 1e3, "pairs_per_sec": p * n / t, "pods_per_sec": p / t}
            if place:
                tp = statistics.mean(place) * 1e-3
                sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
                e["place_kernel_ms"] = tp * 1e3
                e["issue_frac_3op"] = 3.0 * p * n / tp / (sm_count * 128 * measured_peaks()[1] * 1e6)
            extras[name] = e
        except Exception as ex:  # noqa: BLE001
            extras[name] = {"error": repr(ex)}
            lib.kp_profile_enable(0)
    try:
        extras["ingest"] = run_ingest_extra(torch, lib, dev)
    except Exception as ex:  # noqa: BLE001
        extras["ingest"] = {"error": repr(ex)}
    try:
        extras["cache_keys"] = run_cachekey_extra(torch, lib, dev)
    except Exception as ex:  # noqa: BLE001
        extras["cache_keys"] = {"error": repr(ex)}
    return extras


def run_cachekey_extra(torch, lib, dev, n_pods=100_000, n_nodes=1000, cpu_sample=3000):
    """RequestCache keys (md5 of pod prefix + shared node text) for a batch of
    pods in one kernel; hashlib on one core over a sample as the CPU baseline."""
    import ctypes
    import hashlib
    from k8s_llm_scheduler_b200 import cachekey, ingest, synth
    from k8s_llm_scheduler_b200.tables import NodeMetrics, PodSpec
    nodes, pods = synth.config2(n_nodes, n_pods)
    recs = [NodeMetrics(f"node-{i}", float(nodes.cpu_usage_pct[i]), float(nodes.mem_usage_pct[i]), 32.0, 128.0,
                        int(nodes.pod_count[i]), int(nodes.max_pods[i])) for i in range(n_nodes)]
    specs = [PodSpec(f"p{j}", "default", c / 1000.0, m / 2 ** 30) for j, (c, m) in
             enumerate(zip(pods.req_cpu_milli.tolist(), pods.req_mem_bytes.tolist()))]
    text = cachekey.nodes_text(recs)
    t0 = time.perf_counter()
    prefix, off = ingest.pack_strings([cachekey.pod_prefix(p) for p in specs])
    host_s = time.perf_counter() - t0
    d_prefix, d_off = torch.from_numpy(prefix).to(dev), torch.from_numpy(off).to(dev)
    d_text = torch.from_numpy(np.frombuffer(bytearray(text, "utf-8"), dtype=np.uint8)).to(dev)
    out = torch.empty((n_pods, 16), dtype=torch.uint8, device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def launch():
        assert lib.kp_cache_keys(d_prefix.data_ptr(), d_off.data_ptr(), n_pods, d_text.data_ptr(), d_text.numel(),
                                 out.data_ptr(), stream) == 0

    for _ in range(3):
        launch()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    a.record()
    for _ in range(reps):
        launch()
    b.record()
    torch.cuda.synchronize()
    t_dev = a.elapsed_time(b) * 1e-3 / reps
    t0 = time.perf_counter()
    want = [hashlib.md5((cachekey.pod_prefix(p) + text).encode()).hexdigest() for p in specs[:cpu_sample]]
    cpu_s = time.perf_counter() - t0
    got = [row.tobytes().hex() for row in out[:cpu_sample].cpu().numpy()]
    hashed = float(n_pods) * (len(text) + float(off[-1]) / n_pods)
    return {"workload": f"{n_pods} pods x {n_nodes} nodes, node text {len(text)} bytes",
            "device_ms": t_dev * 1e3, "keys_per_sec_device": n_pods / t_dev, "hashed_GB_per_sec": hashed / t_dev / 1e9,
            "host_prefix_format_ms": host_s * 1e3,
            "cpu_baseline": {"value": cpu_sample / cpu_s, "unit": "keys/s", "cores": 1, "kind": "port",
                             "sample": f"first {cpu_sample} pods, hashlib.md5 over the reference's key string"},
            "parity_vs_cpu_sample": got == want}


def run_ingest_extra(torch, lib, dev, n_pods=1_000_000, cpu_sample=40_000):
    """pod request strings -> table columns (kp_ingest_pod_requests).
    Device time with the packed string columns resident in HBM (HBM-bound byte work), the whole
    host call (packing in Python + copies + kernel), and the reference's per-object parsing
    (literal mirror, one core) on a sample."""
    import ctypes
    from k8s_llm_scheduler_b200 import ingest, tables
    from k8s_llm_scheduler_b200._lib import KpPodRequestStrings
    rng = np.random.Generator(np.random.PCG64(77))
    k = rng.integers(1, 4, n_pods)
    cm = rng.integers(1, 8000, int(k.sum()))
    mm = rng.integers(1, 16384, int(k.sum()))
    it = iter(zip(cm.tolist(), mm.tolist()))
    pods = [[{"cpu": f"{c}m", "memory": f"{m}Mi"} for c, m in (next(it) for _ in range(kk))] for kk in k.tolist()]
    t0 = time.perf_counter()
    begin, ct, co, mt, mo, _ = ingest.pack_pod_requests(pods)          # the packer the product path uses
    pack_s = time.perf_counter() - t0
    d = [torch.from_numpy(a).to(dev) for a in (begin, ct, co, mt, mo)]
    outs = [torch.empty(n_pods, dtype=dt, device=dev) for dt in (torch.int64, torch.int64, torch.float64, torch.float64, torch.uint8)]
    st = KpPodRequestStrings(n_pods, *[t.data_ptr() for t in d])
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def launch():
        rc = lib.kp_ingest_pod_requests(ctypes.byref(st), *[ctypes.c_void_p(o.data_ptr()) for o in outs], stream)
        assert rc == 0

    for _ in range(3):
        launch()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    a.record()
    for _ in range(reps):
        launch()
    b.record()
    torch.cuda.synchronize()
    t_dev = a.elapsed_time(b) * 1e-3 / reps
    algo_bytes = sum(t.numel() * t.element_size() for t in d) + sum(o.numel() * o.element_size() for o in outs)
    flagged = float((outs[4] != 0).double().mean().item())   # fp64 sums like 0.1 + 0.2 have no integer milli value
    assert int((outs[4] & 4).max().item()) == 0              # nothing outside the device grammar
    t0 = time.perf_counter()
    res = ingest.ingest_pod_requests(pods, device=dev, exact=False)
    e2e_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    want = tables.PodTable.from_container_requests(pods[:cpu_sample], exact=False)
    cpu_s = time.perf_counter() - t0
    ok = bool(np.array_equal(res.table.req_cpu_milli[:cpu_sample], want.req_cpu_milli) and
              np.array_equal(res.table.req_mem_bytes[:cpu_sample], want.req_mem_bytes))
    hbm_peak, _, _ = measured_peaks()
    return {"workload": f"{n_pods} pods, {int(k.sum())} containers, cpu 'Nm' / memory 'NMi' strings",
            "device_ms": t_dev * 1e3, "pods_per_sec_device": n_pods / t_dev,
            "roofline": {"bound": "hbm", "achieved": algo_bytes / t_dev / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": algo_bytes / t_dev / 1e9 / hbm_peak, "bytes_per_launch": algo_bytes},
            "rows_redone_on_host_frac": flagged, "host_pack_ms": pack_s * 1e3, "e2e_ms": e2e_s * 1e3, "pods_per_sec_e2e": n_pods / e2e_s,
            "cpu_baseline": {"value": cpu_sample / cpu_s, "unit": "pods/s", "cores": 1, "kind": "port",
                             "sample": f"first {cpu_sample} pods through the literal mirror of _convert_pod_to_spec"},
            "parity_vs_cpu_sample": ok}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c4shard", "c5", "c5greedy"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU-baseline sample budget")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--extra", action=argparse.BooleanOptionalAction, default=True)
    ap.add_argument("--no-flush", action="store_true", help="diagnostic: do not flush L2 between timed steps")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
