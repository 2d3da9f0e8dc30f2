"""kp_snapshot_place_noalloc: the snapshot placement for hosts that cannot allocate device
memory when they call us (GPU memory fully owned by another runtime's pool, or a latency
budget that rules out cudaMalloc/cudaFree).  Its scratch is a fixed array inside the
library, so consecutive calls reuse the same bytes: the second call must not see anything
the first one left behind.  Checked bit-for-bit against kp_snapshot_place (which gets a
fresh torch-allocated workspace) and against the CPU oracle, through the C ABI, on calls in
a row.

Collected last on purpose (file name): if it fails under -x every other test has already run.
"""
import ctypes
from ctypes import byref, c_void_p

import numpy as np
import pytest

from k8s_llm_scheduler_b200 import synth
from tests.helpers import assert_same_f64, assert_same_idx, oracle_snapshot

pytestmark = pytest.mark.gpu


def _call_noalloc(lib, placement, dn, rc, rm, strategy, filter_mode):
    """Inputs and outputs are ordinary caller-owned device buffers; no workspace is passed."""
    import torch
    nst, device, n = placement._node_struct(dn)
    pst, p, strat, _ = placement._pod_struct(rc, rm, strategy, device)
    out_idx = torch.full((p,), -7, dtype=torch.int64, device=device)
    out_score = torch.full((p,), -7.0, dtype=torch.float64, device=device)
    stream = torch.cuda.current_stream(device).cuda_stream
    st = lib.kp_snapshot_place_noalloc(byref(nst), byref(pst), strat, filter_mode, out_idx.data_ptr(),
                                       out_score.data_ptr(), c_void_p(stream))
    return st, out_idx, out_score


def test_two_calls_in_a_row_match_the_workspace_entry_point():
    import torch
    from k8s_llm_scheduler_b200 import _lib, placement
    lib = _lib.load()
    # call 1: big table (fills most of the built-in scratch), per-pod strategies
    # call 2: small table, uniform strategy -> must not be affected by what call 1 left behind
    # call 3: the reference's own filter, which uses a different part of the scratch
    cases = [(synth.config5(8000, 60_000, seed=71), "per_pod", 1),
             (synth.config5(700, 5_000, seed=72), 1, 1),
             (synth.config5(3000, 20_000, seed=73), "per_pod", 0),
             (synth.config2(1000, 100_000), 1, 1)]
    for (nodes, pods), strategy, filter_mode in cases:
        dn = placement.nodes_to_device(nodes, "cuda:0")
        rc = torch.as_tensor(pods.req_cpu_milli).cuda()
        rm = torch.as_tensor(pods.req_mem_bytes).cuda()
        strat = torch.as_tensor(pods.strategy_ids).cuda() if strategy == "per_pod" else strategy
        st, got_idx, got_score = _call_noalloc(lib, placement, dn, rc, rm, strat, filter_mode)
        assert st == 0, lib.kp_last_error()
        want_idx, want_score = placement.snapshot_place(dn, rc, rm, strat, filter_mode)
        torch.cuda.synchronize()
        assert torch.equal(got_idx, want_idx)
        assert np.array_equal(got_score.cpu().numpy().view(np.uint64), want_score.cpu().numpy().view(np.uint64))
        assert (got_idx >= -1).all()
        oi, os_ = oracle_snapshot(nodes, pods, strategy if strategy != "per_pod" else 0, filter_mode,
                                  per_pod=strategy == "per_pod")
        assert_same_idx(got_idx.cpu().numpy(), oi)
        assert_same_f64(got_score.cpu().numpy(), os_)


def test_sizes_beyond_the_built_in_scratch_are_rejected():
    import torch
    from k8s_llm_scheduler_b200 import _lib, placement
    lib = _lib.load()
    nodes, pods = synth.config2(8193, 64)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    st, idx, score = _call_noalloc(lib, placement, dn, rc, rm, 0, 1)
    assert st == 5 and b"8193 nodes exceed" in lib.kp_last_error()
    assert (idx == -7).all()                                   # nothing was written
    nodes, pods = synth.config5(100, 65_537)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    sid = torch.as_tensor(pods.strategy_ids).cuda()
    st, idx, _ = _call_noalloc(lib, placement, dn, rc, rm, sid, 1)
    assert st == 5 and b"per-pod strategy" in lib.kp_last_error()
    st, idx, _ = _call_noalloc(lib, placement, dn, rc, rm, 2, 1)      # uniform strategy: no pod limit
    assert st == 0 and (idx >= -1).all()
