"""Out-of-bounds writes: compute-sanitizer is not available on the GPU pool, so the kernels run here with
their workspace and outputs carved out of larger buffers whose surroundings are filled with a pattern;
the pattern must survive.  Sizes are chosen off every tile / chunk boundary, and the workspace is exactly
what kp_*_workspace_bytes reports."""
import ctypes

import numpy as np
import pytest
import torch

from k8s_llm_scheduler_b200 import synth
from tests.helpers import assert_same_idx, oracle_greedy, oracle_snapshot

pytestmark = pytest.mark.gpu
PAD = 4096


class Guarded:
    def __init__(self, nbytes, device="cuda:0"):
        self.nbytes = int(nbytes)
        self.raw = torch.full((PAD + (self.nbytes + 255) // 256 * 256 + PAD,), 0xA5, dtype=torch.uint8, device=device)

    def view(self, dtype=torch.uint8):
        return self.raw[PAD:PAD + self.nbytes].view(dtype)

    def intact(self):
        torch.cuda.synchronize()
        tail = self.raw[PAD + self.nbytes:]
        return bool((self.raw[:PAD] == 0xA5).all()) and bool((tail == 0xA5).all())


@pytest.mark.parametrize("n,p,per_pod", [(700, 3001, False), (700, 3001, True), (1, 1, False), (8192, 777, True),
                                         (9001, 2503, False), (9001, 2503, True), (33_000, 4099, False)])
def test_snapshot_stays_inside_its_buffers(n, p, per_pod):
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = synth.config5(n, p, seed=n + p)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    strat = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else 1
    ws = Guarded(placement.snapshot_workspace_bytes(n, p, per_pod))
    oi, os_ = Guarded(8 * p), Guarded(8 * p)
    for mode in (1, 0):
        idx, score = placement.snapshot_place(dn, rc, rm, strat, mode, workspace=ws.view(),
                                              out=(oi.view(torch.int64), os_.view(torch.float64)))
        assert ws.intact() and oi.intact() and os_.intact(), f"mode {mode}: a kernel wrote outside its buffers"
        want, _ = oracle_snapshot(nodes, pods, 1, mode, per_pod=per_pod)
        assert_same_idx(idx.cpu().numpy(), want)


@pytest.mark.parametrize("n,p,per_pod", [(300, 2049, False), (300, 2049, True), (6000, 2500, True), (21_000, 2100, False),
                                         (50, 100, False)])
def test_greedy_stays_inside_its_workspace(n, p, per_pod):
    from k8s_llm_scheduler_b200 import placement
    nodes, pods = synth.config5(n, p, seed=n + p)
    dn = placement.nodes_to_device(nodes, "cuda:0")
    rc, rm = torch.as_tensor(pods.req_cpu_milli).cuda(), torch.as_tensor(pods.req_mem_bytes).cuda()
    strat = torch.as_tensor(pods.strategy_ids).cuda() if per_pod else 0
    ws = Guarded(placement.greedy_workspace_bytes(n, p))
    out = placement.greedy_place(dn, rc, rm, strat, 1, workspace=ws.view())
    assert ws.intact(), "a greedy kernel wrote outside its workspace"
    want = oracle_greedy(nodes, pods, 0, 1, per_pod=per_pod)
    assert_same_idx(out[0].cpu().numpy(), want[0])


def test_ingest_stays_inside_its_outputs():
    from k8s_llm_scheduler_b200 import _lib
    from k8s_llm_scheduler_b200.ingest import flatten_pod_requests, pack_strings
    lib = _lib.load()
    rng = np.random.Generator(np.random.PCG64(5))
    pods = [[{"cpu": f"{int(rng.integers(0, 9000))}m", "memory": f"{int(rng.integers(0, 9000))}Mi"}
             for _ in range(int(rng.integers(0, 4)))] for _ in range(1001)]
    begin, cpu, mem, _ = flatten_pod_requests(pods)
    (ct, co), (mt, mo) = pack_strings(cpu), pack_strings(mem)
    dev = [torch.from_numpy(a).cuda() for a in (begin, ct, co, mt, mo)]
    n = len(pods)
    outs = [Guarded(8 * n), Guarded(8 * n), Guarded(8 * n), Guarded(8 * n), Guarded(n)]
    st = _lib.KpPodRequestStrings(n, *[t.data_ptr() for t in dev])
    rc = lib.kp_ingest_pod_requests(ctypes.byref(st), *[g.view().data_ptr() for g in outs], None)
    assert rc == 0
    assert all(g.intact() for g in outs), "the ingest kernel wrote outside its outputs"
    assert int(outs[4].view().max().item()) < 4
