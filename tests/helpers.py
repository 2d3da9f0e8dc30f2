"""Shared helpers for the parity tests: run the CUDA path through the C ABI, run the
oracle on the same tables, compare bit-for-bit."""
import struct

import numpy as np

from k8s_llm_scheduler_b200.tables import NodeTable, PodTable


def bits(x) -> str:
    return struct.pack(">d", float(x)).hex()


def from_bits(h: str) -> float:
    return struct.unpack(">d", bytes.fromhex(h))[0]


def assert_same_f64(got: np.ndarray, want: np.ndarray, what: str = "score"):
    """Bit-exact fp64 comparison (stricter than the 1e-9 relative the north star allows)."""
    g = np.ascontiguousarray(got, dtype=np.float64).view(np.uint64)
    w = np.ascontiguousarray(want, dtype=np.float64).view(np.uint64)
    bad = np.nonzero(g != w)[0]
    assert bad.size == 0, f"{what}: {bad.size} mismatches, first at {bad[0]}: " \
                          f"got {got[bad[0]]!r} want {want[bad[0]]!r}"


def assert_same_idx(got: np.ndarray, want: np.ndarray, what: str = "node index"):
    got, want = np.asarray(got), np.asarray(want)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, f"{what}: {bad.size} mismatches, first at pod {bad[0]}: " \
                          f"got {got[bad[0]]} want {want[bad[0]]}"


def device_snapshot(nodes: NodeTable, pods: PodTable, strategy=0, filter_mode=1, per_pod=False):
    """CUDA snapshot path via kp_snapshot_place (device pointers, torch-owned memory)."""
    import torch
    from k8s_llm_scheduler_b200 import placement
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc = torch.as_tensor(pods.req_cpu_milli).cuda()
    rm = torch.as_tensor(pods.req_mem_bytes).cuda()
    strat = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else strategy
    idx, score = placement.snapshot_place(dn, rc, rm, strat, filter_mode)
    torch.cuda.synchronize()
    return idx.cpu().numpy(), score.cpu().numpy()


def device_greedy(nodes: NodeTable, pods: PodTable, strategy=0, filter_mode=1, per_pod=False):
    import torch
    from k8s_llm_scheduler_b200 import placement
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc = torch.as_tensor(pods.req_cpu_milli).cuda()
    rm = torch.as_tensor(pods.req_mem_bytes).cuda()
    strat = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else strategy
    out = placement.greedy_place(dn, rc, rm, strat, filter_mode)
    torch.cuda.synchronize()
    return tuple(o.cpu().numpy() for o in out)


def oracle_snapshot(nodes: NodeTable, pods: PodTable, strategy=0, filter_mode=1, per_pod=False, threads=8):
    from oracle import c_oracle
    return c_oracle.snapshot(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes, strategy=strategy,
                             strategy_ids=pods.strategy_ids if per_pod else None,
                             filter_mode=filter_mode, threads=threads)


def oracle_greedy(nodes: NodeTable, pods: PodTable, strategy=0, filter_mode=1, per_pod=False):
    from oracle import c_oracle
    return c_oracle.greedy(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes, strategy=strategy,
                           strategy_ids=pods.strategy_ids if per_pod else None, filter_mode=filter_mode)


def state_after(nodes: NodeTable, idx_prefix: np.ndarray) -> NodeTable:
    """Node table as the reference's loop would see it after the binds in ``idx_prefix`` (-1 = pod not
    bound): pod_count += binds (scheduler.py:598 -> :144-147), both usages re-derived per :150-151 for
    every node that received a pod, untouched nodes keep the caller's values.  Greedy placement is Markov
    in this state, so an oracle restarted from it must continue exactly like the full run."""
    from k8s_llm_scheduler_b200 import synth
    placed = idx_prefix[idx_prefix >= 0]
    add = np.bincount(placed, minlength=len(nodes)).astype(np.int64)
    pc = nodes.pod_count + add
    u = synth.usage_a3(pc, nodes.max_pods)
    touched = add > 0
    return NodeTable.from_columns(alloc_cpu_milli=nodes.alloc_cpu_milli, alloc_mem_bytes=nodes.alloc_mem_bytes,
                                  pod_count=pc, max_pods=nodes.max_pods, ready=nodes.ready,
                                  cpu_usage_pct=np.where(touched, u, nodes.cpu_usage_pct),
                                  mem_usage_pct=np.where(touched, u, nodes.mem_usage_pct))


def check_greedy_from_checkpoint(nodes: NodeTable, pods: PodTable, got_idx, got_score, k: int, count: int,
                                 strategy=0, filter_mode=1, per_pod=False, what=""):
    """Restart the C oracle from the state the device's own first k decisions imply and compare the next
    ``count`` decisions (indices and score bits)."""
    hi = min(len(pods), k + count)
    state = state_after(nodes, np.asarray(got_idx[:k]))
    part = PodTable.from_columns(pods.req_cpu_milli[k:hi], pods.req_mem_bytes[k:hi],
                                 pods.strategy_ids[k:hi] if per_pod else None)
    o = oracle_greedy(state, part, strategy, filter_mode, per_pod=per_pod)
    assert_same_idx(got_idx[k:hi], o[0], f"{what} pods {k}..{hi}")
    assert_same_f64(got_score[k:hi], o[1], f"{what} scores of pods {k}..{hi}")
    return o


COND = {
    "ready": [{"type": "Ready", "status": "True", "reason": "x"}],
    "false": [{"type": "Ready", "status": "False", "reason": "x"}],
    "unknown": [{"type": "Ready", "status": "Unknown", "reason": "x"}],
    "none": [],
    "other_only": [{"type": "MemoryPressure", "status": "True", "reason": "x"}],
    "ready_second": [{"type": "DiskPressure", "status": "False", "reason": "x"},
                     {"type": "Ready", "status": "True", "reason": "x"}],
}
