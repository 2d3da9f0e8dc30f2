"""Host-side multi-rank logic (sharding, table broadcast, arg-max reduce) under gloo with
world_size 2 on CPU.  The local placement op is replaced by the CPU oracle (test
infrastructure), so what is exercised here is k8s_llm_scheduler_b200.dist itself."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200 import dist as kdist
from k8s_llm_scheduler_b200.tables import NODE_COLUMNS


def oracle_place(nodes, rc, rm, strategy, filter_mode):
    """CPU stand-in with the signature of placement.snapshot_place."""
    from oracle import c_oracle
    cols = {k: nodes[k].numpy() for k in NODE_COLUMNS}
    per_pod = isinstance(strategy, torch.Tensor)
    idx, score = c_oracle.snapshot(cols, rc.numpy(), rm.numpy(), strategy=0 if per_pod else strategy,
                                   strategy_ids=strategy.numpy() if per_pod else None, filter_mode=filter_mode)
    return torch.from_numpy(idx), torch.from_numpy(score)


def node_tensors(nodes):
    return {k: torch.from_numpy(np.ascontiguousarray(getattr(nodes, k))) for k in NODE_COLUMNS}


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nodes, pods = synth.config5(700, 3001, seed=77)
        full = node_tensors(nodes)
        rc, rm = torch.from_numpy(pods.req_cpu_milli), torch.from_numpy(pods.req_mem_bytes)
        sid = torch.from_numpy(pods.strategy_ids)
        want_idx, want_score = oracle_place(full, rc, rm, sid, 1)

        # pod-sharded: only rank 0 holds the real table; the others pass zero-filled buffers
        mine = full if rank == 0 else {k: torch.zeros_like(v) for k, v in full.items()}
        gi, gs = kdist.snapshot_pod_sharded(mine, rc, rm, sid, 1, place_fn=oracle_place, gather=True)
        assert torch.equal(gi, want_idx) and torch.equal(gs.view(torch.int64), want_score.view(torch.int64))
        lo, hi, li, ls = kdist.snapshot_pod_sharded(mine, rc, rm, 0, 1, place_fn=oracle_place)
        w2i, w2s = oracle_place(full, rc, rm, 0, 1)
        assert (lo, hi) == kdist.shard_bounds(3001, rank, world)
        assert torch.equal(li, w2i[lo:hi]) and torch.equal(ls.view(torch.int64), w2s[lo:hi].view(torch.int64))

        # node-sharded: contiguous node ranges, arg-max reduce under (score desc, index asc)
        nlo, nhi = kdist.shard_bounds(700, rank, world)
        for strat, want in ((0, oracle_place(full, rc, rm, 0, 1)), (1, oracle_place(full, rc, rm, 1, 0)),
                            (2, oracle_place(full, rc, rm, 2, 1))):
            fm = 0 if strat == 1 else 1
            ni, ns = kdist.snapshot_node_sharded(kdist.slice_nodes(full, nlo, nhi), nlo, rc, rm, strat, fm,
                                                 place_fn=oracle_place)
            assert torch.equal(ni, want[0]), f"strategy {strat}"
            assert torch.equal(ns.view(torch.int64), want[1].view(torch.int64))
        # every node identical: ties everywhere, the first eligible node of the whole list wins
        eq = node_tensors(synth.all_equal_nodes(700))
        eq["ready"][:400] = 0                      # all of rank 0's nodes and some of rank 1's are not Ready
        ni, _ = kdist.snapshot_node_sharded(kdist.slice_nodes(eq, nlo, nhi), nlo, rc[:50], rm[:50], 0, 0,
                                            place_fn=oracle_place)
        assert (ni == 400).all()
        out[rank] = 1
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_snapshot_sharding():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: 1, 1: 1}


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 100, 3001):
        for w in (1, 2, 3, 8):
            spans = [kdist.shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_single_process_paths_need_no_process_group():
    nodes, pods = synth.config5(120, 300, seed=5)
    full = node_tensors(nodes)
    rc, rm = torch.from_numpy(pods.req_cpu_milli), torch.from_numpy(pods.req_mem_bytes)
    want = oracle_place(full, rc, rm, 0, 1)
    gi, gs = kdist.snapshot_pod_sharded(full, rc, rm, 0, 1, place_fn=oracle_place, gather=True)
    assert torch.equal(gi, want[0])
    ni, ns = kdist.snapshot_node_sharded(full, 0, rc, rm, 0, 1, place_fn=oracle_place)
    assert torch.equal(ni, want[0]) and torch.equal(ns.view(torch.int64), want[1].view(torch.int64))
