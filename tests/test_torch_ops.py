"""The PyTorch-extension form of the ops (csrc/kp_torch_ext.cpp): TORCH_LIBRARY(k8s_place, ...) over the C ABI.
CPU part: the extension builds, loads and registers its schemas (no compute without a GPU).  GPU part: the ops,
called through torch.ops directly, give the oracle's answers; there is no CPU kernel to fall back to."""
import numpy as np
import pytest
import torch

from k8s_llm_scheduler_b200 import synth
from k8s_llm_scheduler_b200.build import build_torch_ext
from k8s_llm_scheduler_b200.tables import NODE_COLUMNS


@pytest.fixture(scope="module")
def ops():
    build_torch_ext()
    from k8s_llm_scheduler_b200 import placement
    return placement.ops()


def test_extension_registers_its_operators(ops):
    for name in ("snapshot_place", "snapshot_place_out", "snapshot_place_pruned", "greedy_place", "node_scores"):
        schema = getattr(ops, name).default._schema
        assert schema.name == f"k8s_place::{name}"
    args = [a.name for a in ops.snapshot_place.default._schema.arguments]
    assert args[:7] == list(NODE_COLUMNS) and args[7:] == ["req_cpu_milli", "req_mem_bytes", "strategy_ids", "strategy", "filter_mode"]
    assert ops.snapshot_workspace_bytes(1000, 100_000, False) > 0 and ops.greedy_workspace_bytes(5000, 1_000_000) > 0


def test_no_cpu_kernel_is_registered(ops):
    nodes, pods = synth.config2(4, 4)
    cols = [torch.as_tensor(getattr(nodes, k)) for k in NODE_COLUMNS]
    with pytest.raises((NotImplementedError, RuntimeError)):
        ops.snapshot_place(*cols, torch.as_tensor(pods.req_cpu_milli), torch.as_tensor(pods.req_mem_bytes), None, 0, 1)


@pytest.mark.gpu
def test_ops_match_the_oracle(ops):
    from tests.helpers import assert_same_f64, assert_same_idx, oracle_greedy, oracle_snapshot
    nodes, pods = synth.config5(3000, 9000, seed=11)
    cols = [torch.as_tensor(getattr(nodes, k)).cuda() for k in NODE_COLUMNS]
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    sid = torch.as_tensor(pods.strategy_ids).cuda()
    idx, score = ops.snapshot_place(*cols, rc, rm, sid, 0, 1)
    oi, os_ = oracle_snapshot(nodes, pods, 0, 1, per_pod=True)
    assert_same_idx(idx.cpu().numpy(), oi)
    assert_same_f64(score.cpu().numpy(), os_)
    out_idx, out_score = torch.empty_like(idx), torch.empty_like(score)
    ws = torch.empty(ops.snapshot_workspace_bytes(3000, 9000, False), dtype=torch.uint8, device="cuda")
    ops.snapshot_place_out(*cols, rc, rm, None, 1, 1, out_idx, out_score, ws)
    oi, os_ = oracle_snapshot(nodes, pods, 1, 1)
    assert_same_idx(out_idx.cpu().numpy(), oi)
    assert_same_f64(out_score.cpu().numpy(), os_)
    pi, ps = ops.snapshot_place_pruned(*cols, rc, rm, None, 1, 1, 16)           # same answers with the probe in front
    assert_same_idx(pi.cpu().numpy(), oi)
    assert_same_f64(ps.cpu().numpy(), os_)
    g = ops.greedy_place(*cols, rc, rm, None, 0, 1, None)
    o = oracle_greedy(nodes, pods, 0, 1)
    assert_same_idx(g[0].cpu().numpy(), o[0])
    assert np.array_equal(g[2].cpu().numpy(), o[2])
    sc, fc, fm, el = ops.node_scores(*cols, 0, 1)
    assert sc.dtype == torch.float64 and el.dtype == torch.uint8 and int(el.sum()) > 0
    # wrong dtype / shape / device are rejected with a message, not a crash
    with pytest.raises(RuntimeError, match="dtype"):
        ops.snapshot_place(*cols, rc.to(torch.int32), rm, None, 0, 1)
    with pytest.raises(RuntimeError, match="shape"):
        ops.snapshot_place(*cols, rc, rm[:-1], None, 0, 1)
    with pytest.raises(RuntimeError, match="filter_mode"):
        ops.snapshot_place(*cols, rc, rm, None, 0, 7)
