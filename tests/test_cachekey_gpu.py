"""kp_cache_keys against the reference's RequestCache._generate_key (scheduler.py:265-271):
golden keys produced by the unmodified reference (tests/golden/cache_keys.json) and hashlib on
larger seeded batches (message lengths across every MD5 padding case)."""
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest

from k8s_llm_scheduler_b200.tables import NodeMetrics, PodSpec
from tests.helpers import from_bits

pytestmark = pytest.mark.gpu
GOLD = json.loads((Path(__file__).parent / "golden" / "cache_keys.json").read_text())


def test_golden_keys_of_the_reference():
    from k8s_llm_scheduler_b200.cachekey import cache_keys
    assert len(GOLD) >= 5
    for case in GOLD:
        nodes = [NodeMetrics(n["name"], from_bits(n["cu_bits"]), from_bits(n["cu_bits"]), 8.0, 16.0, n["pc"], n["mp"])
                 for n in case["nodes"]]
        pods = [PodSpec(f"p{j}", "default", from_bits(p["cpu_bits"]), from_bits(p["mem_bits"]), priority=p["priority"])
                for j, p in enumerate(case["pods"])]
        assert cache_keys(pods, nodes) == case["keys"], len(nodes)


@pytest.mark.parametrize("n_nodes", [0, 1, 2, 3, 4, 5, 7, 1000])
def test_matches_hashlib_over_all_padding_cases(n_nodes):
    from k8s_llm_scheduler_b200.cachekey import cache_keys, nodes_text, pod_prefix
    rng = np.random.Generator(np.random.PCG64(n_nodes))
    nodes = [NodeMetrics("n" * int(rng.integers(1, 40)) + str(i), float(rng.random() * 50), float(rng.random() * 50),
                         8.0, 16.0, 1, 110) for i in range(n_nodes)]
    pods = [PodSpec("p", "default", float(rng.integers(0, 10 ** int(rng.integers(1, 9)))) / 1000.0,
                    float(rng.integers(0, 1 << 20)) / 1024, priority=int(rng.integers(-5, 10 ** int(rng.integers(1, 9)))))
            for _ in range(3000)]
    text = nodes_text(nodes)
    want = [hashlib.md5((pod_prefix(p) + text).encode()).hexdigest() for p in pods]
    assert cache_keys(pods, nodes) == want
    lengths = {(len(pod_prefix(p)) + len(text)) % 64 for p in pods}
    if n_nodes in (0, 1, 2, 3):
        assert len(lengths) >= 8              # several message lengths modulo the block size in one batch
    assert cache_keys([], nodes) == []
