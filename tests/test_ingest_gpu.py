"""Parity of the ingest kernels (kp_ingest_pod_requests / kp_ingest_nodes) with the reference's
parsers: golden vectors produced by the unmodified reference (tests/golden/ingest.json), the
literal host mirror on large seeded batches, and the invariants the status flags promise."""
import json
from pathlib import Path

import numpy as np
import pytest

from k8s_llm_scheduler_b200 import tables
from tests.helpers import assert_same_f64, assert_same_idx, bits, from_bits

pytestmark = pytest.mark.gpu

GOLD = json.loads((Path(__file__).parent / "golden" / "ingest.json").read_text())
HOST, CPU_INEXACT, MEM_INEXACT = 4, 1, 2


def test_pod_golden_vectors_raw_device_results():
    from k8s_llm_scheduler_b200.ingest import ingest_pod_requests
    recs = GOLD["pods"]
    raw = ingest_pod_requests([r["containers"] for r in recs], redo=False)
    n_dev = 0
    for j, rec in enumerate(recs):
        st = int(raw.status[j])
        if rec["error"]:
            assert st & HOST, f"device accepted what the reference rejects: {rec['containers']}"
            continue
        if st & HOST:
            continue                                     # deferred to the host: checked below
        n_dev += 1
        assert bits(raw.cpu_request[j]) == rec["cpu_bits"], rec["containers"]
        assert bits(raw.memory_request[j]) == rec["mem_bits"], rec["containers"]
        if not st & CPU_INEXACT:
            assert bits(int(raw.table.req_cpu_milli[j]) / 1000.0) == rec["cpu_bits"]
        if not st & MEM_INEXACT:
            assert bits(int(raw.table.req_mem_bytes[j]) / 1024 / 1024 / 1024) == rec["mem_bits"]
    assert n_dev >= 250                                  # the device, not the host, did most rows


def test_pod_golden_vectors_with_host_redo():
    from k8s_llm_scheduler_b200.ingest import ingest_pod_requests
    def in_wire_range(r):                                # the int64 columns hold |value| < 2^53
        c, m = from_bits(r["cpu_bits"]), from_bits(r["mem_bits"])
        return abs(c * 1000) < 2 ** 53 and abs(m * 2 ** 30) < 2 ** 53

    parsed = [r for r in GOLD["pods"] if not r["error"]]
    good = [r for r in parsed if in_wire_range(r)]
    assert len(good) >= 300
    for r in [r for r in parsed if not in_wire_range(r)][:10]:
        with pytest.raises(ValueError):                  # nan / inf / 1e19 cores: no table row for that
            ingest_pod_requests([r["containers"]], exact=False)
    out = ingest_pod_requests([r["containers"] for r in good], exact=False)
    assert_same_f64(out.cpu_request, np.array([from_bits(r["cpu_bits"]) for r in good]), "cpu_request")
    assert_same_f64(out.memory_request, np.array([from_bits(r["mem_bits"]) for r in good]), "memory_request")
    want = tables.PodTable.from_container_requests([r["containers"] for r in good], exact=False)
    assert_same_idx(out.table.req_cpu_milli, want.req_cpu_milli, "req_cpu_milli")
    assert_same_idx(out.table.req_mem_bytes, want.req_mem_bytes, "req_mem_bytes")
    bad = [r for r in GOLD["pods"] if r["error"]]
    for r in bad[:10]:
        with pytest.raises(ValueError):
            ingest_pod_requests([good[0]["containers"], r["containers"]])


def test_node_golden_vectors():
    from k8s_llm_scheduler_b200.ingest import ingest_nodes
    recs = GOLD["nodes"]
    rows = [dict(name=r["name"], cpu=r["cpu"], memory=r["memory"], pods=r["pods"], pod_count=r["n_running"])
            for r in recs]
    raw = ingest_nodes(rows, redo=False)
    n_dev = 0
    for i, rec in enumerate(recs):
        st = int(raw.status[i])
        if not rec["ok"]:
            assert st & HOST, f"device accepted what the reference rejects: {rows[i]}"
            continue
        if st & HOST:
            continue
        n_dev += 1
        assert bits(raw.available_cpu[i]) == rec["available_cpu_bits"], rows[i]
        assert bits(raw.available_memory[i]) == rec["available_memory_bits"], rows[i]
        assert int(raw.table.max_pods[i]) == rec["max_pods"]
        assert bits(raw.table.cpu_usage_pct[i]) == rec["cu_bits"]
        assert bits(raw.table.mem_usage_pct[i]) == rec["mu_bits"]
        if not st & CPU_INEXACT:
            assert bits(int(raw.table.alloc_cpu_milli[i]) / 1000.0) == rec["available_cpu_bits"]
        if not st & MEM_INEXACT:
            assert bits(int(raw.table.alloc_mem_bytes[i]) / 1024 / 1024 / 1024) == rec["available_memory_bits"]
    assert n_dev >= 120
    def in_wire_range(r):
        return r["ok"] and abs(r["max_pods"]) < 2 ** 53 and abs(from_bits(r["available_cpu_bits"]) * 1000) < 2 ** 53 \
            and abs(from_bits(r["available_memory_bits"]) * 2 ** 30) < 2 ** 53

    ok_rows = [rows[i] for i, r in enumerate(recs) if in_wire_range(r)]
    out = ingest_nodes(ok_rows, exact=False)
    ok = [r for r in recs if in_wire_range(r)]
    assert len(ok) >= 150
    assert_same_f64(out.available_cpu, np.array([from_bits(r["available_cpu_bits"]) for r in ok]))
    assert_same_f64(out.available_memory, np.array([from_bits(r["available_memory_bits"]) for r in ok]))
    assert_same_f64(out.table.cpu_usage_pct, np.array([from_bits(r["cu_bits"]) for r in ok]))
    assert_same_idx(out.table.max_pods, np.array([r["max_pods"] for r in ok]))


def _random_pods(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    k = rng.integers(0, 4, n)
    cpu_m = rng.integers(0, 64000, (n, 3))
    mem_mi = rng.integers(0, 65536, (n, 3))
    form = rng.integers(0, 6, (n, 3))
    pods = []
    for j in range(n):
        cs = []
        for c in range(int(k[j])):
            f = int(form[j, c])
            cpu = f"{cpu_m[j, c]}m" if f < 3 else (f"{cpu_m[j, c] / 1000:.3f}" if f < 5 else str(cpu_m[j, c] // 1000))
            mem = (f"{mem_mi[j, c]}Mi", f"{mem_mi[j, c] * 1024}Ki", f"{mem_mi[j, c] // 1024}Gi",
                   f"{mem_mi[j, c] / 1024:.4f}Gi", str(mem_mi[j, c] << 20), f"{mem_mi[j, c] / 8:.3f}Mi")[f]
            cs.append({"cpu": cpu, "memory": mem})
        pods.append(cs)
    return pods


def test_large_random_batch_matches_the_host_mirror():
    from k8s_llm_scheduler_b200.ingest import ingest_pod_requests
    pods = _random_pods(200_000, 11)
    raw = ingest_pod_requests(pods, redo=False)
    assert not (raw.status & HOST).any()                 # canonical spellings never need the host
    sample = np.random.Generator(np.random.PCG64(12)).choice(len(pods), 20_000, replace=False)
    want = np.array([tables.pod_requests(pods[j]) for j in sample])
    assert_same_f64(raw.cpu_request[sample], want[:, 0], "cpu_request")
    assert_same_f64(raw.memory_request[sample], want[:, 1], "memory_request")
    exact = raw.status == 0
    assert exact.mean() > 0.5
    assert_same_f64(raw.table.req_cpu_milli[exact] / 1000.0, raw.cpu_request[exact], "milli wire value")
    assert_same_f64(raw.table.req_mem_bytes[exact] / 1024 / 1024 / 1024, raw.memory_request[exact], "byte wire value")
    # flagged-inexact rows really have no wire value
    for j in np.nonzero(raw.status & CPU_INEXACT)[0][:200]:
        with pytest.raises(ValueError):
            tables.cores_to_milli(float(raw.cpu_request[j]))


def test_ingest_feeds_the_placement_kernels():
    """strings -> ingest kernels -> snapshot kernel == literal host tables -> oracle."""
    from k8s_llm_scheduler_b200.ingest import ingest_nodes, ingest_pod_requests
    from tests.helpers import device_snapshot, oracle_snapshot
    rng = np.random.Generator(np.random.PCG64(5))
    shapes = [("4", "16Gi", "30"), ("16", "64Gi", "58"), ("32000m", "134217728Ki", "110"), ("7910m", "32780300Ki", "110")]
    rows = []
    for i in range(300):
        cpu, mem, mp = shapes[int(rng.integers(0, len(shapes)))]
        rows.append(dict(name=f"n{i}", cpu=cpu, memory=mem, pods=mp, pod_count=int(rng.integers(0, int(mp) + 1))))
    pods = [[{"cpu": f"{int(rng.integers(1, 80)) * 50}m", "memory": f"{int(rng.integers(1, 256)) * 64}Mi"}]
            for _ in range(5000)]
    nt, pt = ingest_nodes(rows).table, ingest_pod_requests(pods).table
    want_nodes = tables.NodeTable.from_allocatable(rows)
    for col in tables.NODE_COLUMNS:
        assert np.array_equal(getattr(nt, col), getattr(want_nodes, col)), col
    for strat in (0, 1, 2):
        gi, gs = device_snapshot(nt, pt, strat, 1)
        oi, os_ = oracle_snapshot(want_nodes, tables.PodTable.from_container_requests(pods), strat, 1)
        assert_same_idx(gi, oi)
        assert_same_f64(gs, os_)


def test_ragged_batches_and_unaligned_text():
    """Pods with very many containers, long (zero-padded) spellings and text buffers that start at
    an odd address: same answers as the host mirror."""
    import ctypes
    import torch
    from k8s_llm_scheduler_b200 import _lib
    from k8s_llm_scheduler_b200.ingest import flatten_pod_requests, ingest_pod_requests, pack_strings
    rng = np.random.Generator(np.random.PCG64(3))
    pods = []
    for j in range(3000):
        k = 1500 if j in (7, 1300) else int(rng.integers(0, 4))
        pad = "0" * int(rng.integers(0, 50)) if 600 <= j < 1200 else ""
        pods.append([{"cpu": f"{pad}{int(rng.integers(0, 4000))}m", "memory": f"{pad}{int(rng.integers(0, 4096))}Mi"}
                     for _ in range(k)])
    raw = ingest_pod_requests(pods, redo=False)
    assert not (raw.status & HOST).any()
    want = np.array([tables.pod_requests(p) for p in pods])
    assert_same_f64(raw.cpu_request, want[:, 0], "cpu_request")
    assert_same_f64(raw.memory_request, want[:, 1], "memory_request")
    # the same batch with both text buffers shifted by one byte
    lib = _lib.load()
    begin, cpu, mem, _ = flatten_pod_requests(pods)
    cols = [pack_strings(cpu), pack_strings(mem)]
    dev = []
    for text, off in cols:
        shifted = torch.zeros(text.size + 17, dtype=torch.uint8, device="cuda")
        shifted[1:1 + text.size] = torch.from_numpy(text).cuda()
        dev += [shifted, torch.from_numpy(off).cuda()]
    dbegin = torch.from_numpy(begin).cuda()
    n = len(pods)
    o_cpu, o_mem = (torch.empty(n, dtype=torch.float64, device="cuda") for _ in range(2))
    o_st = torch.empty(n, dtype=torch.uint8, device="cuda")
    st = _lib.KpPodRequestStrings(n, dbegin.data_ptr(), dev[0].data_ptr() + 1, dev[1].data_ptr(),
                                  dev[2].data_ptr() + 1, dev[3].data_ptr())
    rc = lib.kp_ingest_pod_requests(ctypes.byref(st), None, None, o_cpu.data_ptr(), o_mem.data_ptr(), o_st.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    assert_same_f64(o_cpu.cpu().numpy(), want[:, 0], "cpu_request (unaligned text)")
    assert_same_f64(o_mem.cpu().numpy(), want[:, 1], "memory_request (unaligned text)")


def test_empty_and_degenerate_batches():
    from k8s_llm_scheduler_b200.ingest import ingest_nodes, ingest_pod_requests
    out = ingest_pod_requests([])
    assert len(out.table) == 0
    out = ingest_pod_requests([[], [None], [{}]])
    assert out.table.req_cpu_milli.tolist() == [0, 0, 0] and out.table.req_mem_bytes.tolist() == [0, 0, 0]
    assert out.host_rows == 0
    out = ingest_pod_requests([[{"cpu": "", "memory": ""}], [{"cpu": "-0", "memory": "5"}]])
    assert out.table.req_cpu_milli.tolist() == [0, 0] and out.table.req_mem_bytes.tolist() == [0, 0]
    assert bits(out.cpu_request[1]) == bits(0.0)         # 0.0 + float("-0") is +0.0, as in the reference's sum
    assert len(ingest_nodes([]).table) == 0
