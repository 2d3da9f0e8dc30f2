"""CPU tests of the ingest host logic: string packing, CSR flattening, and the literal mirrors of
the reference's parsers (tables.py, oracle/py_oracle.py) against tests/golden/ingest.json, which
the unmodified reference produced (tests/golden/make_golden.py: _convert_pod_to_spec
scheduler.py:731-764, get_node_metrics :121-170)."""
import json
from pathlib import Path

import numpy as np
import pytest

from k8s_llm_scheduler_b200 import tables
from oracle import py_oracle
from tests.helpers import bits

GOLD = json.loads((Path(__file__).parent / "golden" / "ingest.json").read_text())


def _pack():
    # the packer needs torch only for the device hand-over; import lazily so that a missing
    # CUDA library cannot break the pure-host checks
    from k8s_llm_scheduler_b200.ingest import flatten_pod_requests, pack_strings
    return flatten_pod_requests, pack_strings


def test_pack_strings_round_trip():
    _, pack_strings = _pack()
    strs = ["250m", "", "1.5Gi", "٣", "x" * 70, "0"]
    data, off = pack_strings(strs)
    assert off.dtype == np.int64 and off[0] == 0 and len(off) == len(strs) + 1
    got = [bytes(data[off[i]:off[i + 1]]).decode("utf-8") for i in range(len(strs))]
    assert got == strs
    data, off = pack_strings([])
    assert data.size == 0 and off.tolist() == [0]
    data, off = pack_strings(["", ""])
    assert data.size == 0 and off.tolist() == [0, 0, 0]


def test_flatten_follows_the_reference_loop():
    flatten, _ = _pack()
    pods = [[{"cpu": "250m", "memory": "256Mi"}], [], [None, {}, {"cpu": "1"}], [{"memory": "1Gi"}, {"cpu": 2}]]
    begin, cpu, mem, odd = flatten(pods)
    assert begin.tolist() == [0, 1, 1, 2, 4]
    assert cpu == ["250m", "1", "0", "0"] and mem == ["256Mi", "0", "1Gi", "0"]
    assert odd == [3]


@pytest.mark.parametrize("mirror", [tables.pod_requests, py_oracle.pod_requests])
def test_pod_mirrors_match_the_reference(mirror):
    assert len(GOLD["pods"]) >= 400
    for rec in GOLD["pods"]:
        if rec["error"]:
            with pytest.raises(ValueError):
                mirror(rec["containers"])
        else:
            c, m = mirror(rec["containers"])
            assert (bits(c), bits(m)) == (rec["cpu_bits"], rec["mem_bits"]), rec["containers"]


def test_node_mirrors_match_the_reference():
    seen_ok = 0
    for rec in GOLD["nodes"]:
        def row():
            mp = int(rec["pods"])
            return (tables.parse_cpu(rec["cpu"]), tables.parse_memory(rec["memory"]), mp,
                    tables.usage_percent(rec["n_running"], mp))
        if not rec["ok"]:
            with pytest.raises(ValueError):
                row()
            continue
        seen_ok += 1
        cpu, mem, mp, u = row()
        assert bits(cpu) == rec["available_cpu_bits"] and bits(mem) == rec["available_memory_bits"]
        assert mp == rec["max_pods"] and rec["pod_count"] == rec["n_running"]
        assert bits(u) == rec["cu_bits"] == rec["mu_bits"]
        o = py_oracle.node_row(rec["name"], rec["cpu"], rec["memory"], rec["pods"], rec["n_running"])
        assert bits(o.available_cpu) == rec["available_cpu_bits"]
        assert bits(o.available_memory) == rec["available_memory_bits"]
        assert bits(o.cpu_usage_percent) == rec["cu_bits"]
    assert seen_ok >= 150


def test_inexact_rows_are_rounded_or_refused_like_the_host_tables():
    """Rows the device parsed but flagged as having no exact wire value: same outcome as tables.py."""
    from k8s_llm_scheduler_b200.ingest import ROW_HOST, _round_inexact
    values = np.array([0.25, 0.1 + 0.2, 1e13, 0.30000000000000004, -0.0, 7.0])
    status = np.array([0, 1, 1 | ROW_HOST, 1, 1, 0], dtype=np.uint8)
    wire = np.array([250, -1, -1, -1, -1, 7000], dtype=np.int64)
    _round_inexact(status, 1, values, 1000.0, wire, tables.cores_to_milli, exact=False)
    assert wire.tolist() == [250, 300, -1, 300, 0, 7000]          # row 2 is the host's, untouched here
    with pytest.raises(ValueError):
        _round_inexact(status, 1, values, 1000.0, wire.copy(), tables.cores_to_milli, exact=True)
    big = np.array([1e16])
    with pytest.raises(ValueError):                                 # outside the int64 wire range even when rounding
        _round_inexact(np.array([1], dtype=np.uint8), 1, big, 1000.0, np.zeros(1, np.int64), tables.cores_to_milli, False)
    only_zero = np.array([-0.0])
    w = np.array([-1], dtype=np.int64)
    _round_inexact(np.array([1], dtype=np.uint8), 1, only_zero, 1000.0, w, tables.cores_to_milli, exact=True)
    assert w.tolist() == [0]                                        # -0.0 == 0: fine for tables.py, fine here


def test_c_packer_equals_the_python_packer():
    """csrc/kp_pyobj.c pack_pod_requests == flatten_pod_requests + pack_strings (non-ASCII, missing keys,
    containers without requests, non-string quantities)."""
    from k8s_llm_scheduler_b200 import tables as T
    from k8s_llm_scheduler_b200.build import build_pyobj
    from k8s_llm_scheduler_b200.ingest import flatten_pod_requests, pack_pod_requests, pack_strings
    build_pyobj()
    T._WALKER[0] = False
    assert T._record_walker() is not None
    pods = [rec["containers"] for rec in GOLD["pods"]] + [[{"cpu": "٣", "memory": "1Gi"}], [], [None, {}], [{"memory": 5}]]
    begin, ct, co, mt, mo, odd = pack_pod_requests(pods)
    b2, cpu, mem, odd2 = flatten_pod_requests(pods)
    (ct2, co2), (mt2, mo2) = pack_strings(cpu), pack_strings(mem)
    assert np.array_equal(begin, b2) and odd == odd2
    assert len(co) == len(co2) == len(mo) == len(mo2) == int(begin[-1]) + 1
    clean = [j for j in range(len(pods)) if j not in set(odd)]          # odd pods carry placeholders on both sides
    for j in clean:
        for c in range(int(begin[j]), int(begin[j + 1])):
            assert bytes(ct[co[c]:co[c + 1]]) == bytes(ct2[co2[c]:co2[c + 1]])
            assert bytes(mt[mo[c]:mo[c + 1]]) == bytes(mt2[mo2[c]:mo2[c + 1]])
