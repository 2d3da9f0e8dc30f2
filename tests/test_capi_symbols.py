"""The C-ABI library builds, loads, and exports every symbol include/*.h declares.
No compute calls here: this runs without a GPU."""
import ctypes
import re
import subprocess
from pathlib import Path

import pytest

from k8s_llm_scheduler_b200 import _lib
from k8s_llm_scheduler_b200.build import LIB_PATH, build_library

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    build_library()
    return _lib.load()


def declared_functions():
    text = (ROOT / "include" / "k8s_place_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kp_[a-z_0-9]+)\s*\(", text)))


def test_header_functions_all_bound_and_exported(lib):
    declared = declared_functions()
    assert len(declared) >= 15
    assert set(declared) == set(_lib.SYMBOLS), "binding table and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", str(LIB_PATH)], capture_output=True, text=True,
                         check=True).stdout
    exported = set(re.findall(r"\bT (kp_[a-z_0-9]+)", out))
    missing = set(declared) - exported
    assert not missing, f"not exported: {missing}"
    for name in declared:
        assert getattr(lib, name) is not None


def test_library_is_sm100a_only(lib):
    out = subprocess.run(["cuobjdump", "-lelf", str(LIB_PATH)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_version_and_status_strings(lib):
    assert lib.kp_abi_version() == 1
    assert lib.kp_status_string(0) == b"ok"
    for code in range(1, 6):
        assert lib.kp_status_string(code) not in (b"ok", b"unknown status")
    assert lib.kp_status_string(99) == b"unknown status"


def test_workspace_sizes_monotone(lib):
    a = lib.kp_snapshot_workspace_bytes(1000, 100000, 0)
    b = lib.kp_snapshot_workspace_bytes(50000, 10_000_000, 0)
    c = lib.kp_snapshot_workspace_bytes(50000, 10_000_000, 1)
    assert 0 < a < b < c
    assert c - b >= 4 * 10_000_000          # the per-pod-strategy permutation
    assert b < 64 << 20                      # node plan only: no P x N matrix, no per-pod scratch
    assert lib.kp_snapshot_workspace_bytes(-1, 5, 0) == 0
    assert lib.kp_greedy_workspace_bytes(5000, 1_000_000) > 0


def test_argument_validation_needs_no_gpu(lib):
    nt, pt = _lib.KpNodeTable(), _lib.KpPodTable()
    nt.n_nodes, pt.n_pods = -1, 0
    st = lib.kp_snapshot_place(ctypes.byref(nt), ctypes.byref(pt), 0, 1, None, None, None, 0, None)
    assert st == 1 and b"negative" in lib.kp_last_error()
    nt.n_nodes = 4                       # NULL columns
    st = lib.kp_greedy_place(ctypes.byref(nt), ctypes.byref(pt), 0, 1, None, None, None, None, None,
                             None, 0, None)
    assert st == 1 and b"NULL" in lib.kp_last_error()
    nt.n_nodes = 0
    assert lib.kp_snapshot_place(ctypes.byref(nt), ctypes.byref(pt), 0, 7, None, None, None, 0, None) == 1
    assert lib.kp_snapshot_place(None, None, 0, 1, None, None, None, 0, None) == 1


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load(tmp_path / "libk8s_place_b200.so")


def test_rebuild_is_decided_by_content_not_by_file_times(lib):
    """After the tree has been copied to another machine file times mean nothing: a library whose stored source hash
    matches is current whatever the mtimes say (a time-based check once made all ranks of a torchrun job rebuild, and
    overwrite, the library at the same moment)."""
    import os
    import time
    from k8s_llm_scheduler_b200 import build as B
    assert B._current(LIB_PATH, B._library_digest())
    src = B.CSRC / "kp_api.cu"
    before = LIB_PATH.stat().st_mtime_ns
    os.utime(src, None)                                   # source now "newer" than the library
    t = time.perf_counter()
    assert build_library() == LIB_PATH
    assert LIB_PATH.stat().st_mtime_ns == before and time.perf_counter() - t < 5.0      # no compiler run
    stamp = LIB_PATH.with_name(LIB_PATH.name + ".srchash")
    good = stamp.read_text()
    try:
        stamp.write_text("0" * 64 + "\n")
        assert not B._current(LIB_PATH, B._library_digest())
    finally:
        stamp.write_text(good)
