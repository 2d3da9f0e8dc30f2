"""The C ABI from a plain C program (examples/c_abi_example.c): compiled with gcc against
include/k8s_place_b200.h and the shared library only, run on the GPU, compared with the oracle."""
import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

from k8s_llm_scheduler_b200.build import LIB_PATH, build_library
from k8s_llm_scheduler_b200.tables import NodeTable, PodTable
from tests.helpers import bits, oracle_greedy, oracle_snapshot

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def test_plain_c_caller_matches_the_oracle(tmp_path):
    if shutil.which("gcc") is None:
        pytest.skip("no gcc on this machine")
    build_library()
    exe = tmp_path / "c_abi_example"
    subprocess.run(["gcc", "-O2", "-Wall", "-Werror", "-I", str(ROOT / "include"), str(ROOT / "examples" / "c_abi_example.c"),
                    "-L", str(LIB_PATH.parent), "-lk8s_place_b200", f"-Wl,-rpath,{LIB_PATH.parent}", "-o", str(exe)],
                   check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True, timeout=120).stdout.splitlines()

    pc = np.array([3, 20, 55, 12, 100, 8])
    mp = np.array([30, 58, 110, 110, 250, 110])
    usage = (pc / mp) * 50
    nodes = NodeTable.from_columns(alloc_cpu_milli=[4000, 16000, 32000, 32000, 96000, 8000],
                                   alloc_mem_bytes=[16 << 30, 64 << 30, 128 << 30, 128 << 30, 768 << 30, 15 << 30],
                                   pod_count=pc, max_pods=mp, ready=[1, 1, 1, 0, 1, 1],
                                   cpu_usage_pct=usage, mem_usage_pct=usage.copy())
    pods = PodTable.from_columns([250, 500, 100, 20000, 64000, 3000, 1, 99000],
                                 [256 << 20, 512 << 20, 128 << 20, 40 << 30, 500 << 30, 10 << 30, 1, 1 << 30])
    want = []
    for sid in range(3):
        oi, os_ = oracle_snapshot(nodes, pods, sid, 1)
        want += [f"snapshot strategy={sid} pod={j} node={oi[j]} score={bits(os_[j])}" for j in range(len(pods))]
        want += [f"decide strategy={sid} pod={j} node={oi[j]} score={bits(os_[j])}" for j in range(len(pods))]
        gi, gs, gpc, gcu, _ = oracle_greedy(nodes, pods, sid, 1)
        want += [f"greedy strategy={sid} pod={j} node={gi[j]} score={bits(gs[j])}" for j in range(len(pods))]
        want += [f"greedy strategy={sid} final node={i} pod_count={gpc[i]} cpu_usage={bits(gcu[i])}" for i in range(len(nodes))]
    want.append("bad filter -> status=1 (invalid argument)")

    def canon(line):                      # the C side prints doubles as %a: compare their bits
        parts = line.split()
        for k, part in enumerate(parts):
            for key in ("score=", "cpu_usage="):
                if part.startswith(key):
                    parts[k] = key + bits(float.fromhex(part[len(key):]))
        return " ".join(parts)

    assert [canon(line) for line in out] == want
