"""Seeded synthetic clusters for the BASELINE.json configs.

Shared byte-for-byte by the tests, bench.py and the oracle so that every arm sees the
same tables.  numpy only (``numpy.random.Generator(PCG64(seed))``).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .tables import NodeTable, PodTable

GIB = 1 << 30
MIB = 1 << 20
HOMOGENEOUS_SHAPE = (32000, 128 * GIB, 110)
SHAPES = [(4000, 16 * GIB, 30), (16000, 64 * GIB, 58), (32000, 128 * GIB, 110), (96000, 768 * GIB, 250)]


def usage_a3(pod_count: np.ndarray, max_pods: np.ndarray) -> np.ndarray:
    """(pod_count / max_pods) * 50, or 0 where max_pods == 0 (scheduler.py:150-151).
    IEEE division then multiplication in fp64 -- the same two roundings as CPython."""
    pc = pod_count.astype(np.float64)
    mp = max_pods.astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        u = (pc / mp) * 50.0
    return np.where(max_pods > 0, u, 0.0)


def homogeneous_nodes(n: int, rng: np.random.Generator, max_count: int = 110) -> NodeTable:
    cpu, mem, mp = HOMOGENEOUS_SHAPE
    pc = rng.integers(0, max_count + 1, n, dtype=np.int64)
    mpa = np.full(n, mp, dtype=np.int64)
    u = usage_a3(pc, mpa)
    return NodeTable.from_columns(alloc_cpu_milli=np.full(n, cpu), alloc_mem_bytes=np.full(n, mem),
                                  pod_count=pc, max_pods=mpa, ready=np.ones(n, np.uint8),
                                  cpu_usage_pct=u, mem_usage_pct=u.copy())


def uniform_pods(p: int, rng: np.random.Generator) -> PodTable:
    cpu = rng.integers(50, 4001, p, dtype=np.int64)
    mem = rng.integers(64, 16 * 1024 + 1, p, dtype=np.int64) * MIB
    return PodTable.from_columns(cpu, mem)


def config1() -> Tuple[NodeTable, PodTable]:
    """3-node minikube cluster x ai-test-pods.yaml."""
    rows = [dict(name="minikube", cpu="8", memory="16374576Ki", pods="110", pod_count=8),
            dict(name="minikube-m02", cpu="8", memory="16374576Ki", pods="110", pod_count=2),
            dict(name="minikube-m03", cpu="8", memory="16374576Ki", pods="110", pod_count=2)]
    nodes = NodeTable.from_allocatable(rows)
    pods = PodTable.from_container_requests([[{"cpu": "250m", "memory": "256Mi"}],
                                             [{"cpu": "500m", "memory": "512Mi"}],
                                             [{"cpu": "100m", "memory": "128Mi"}]])
    return nodes, pods


def config2(n: int = 1000, p: int = 100_000, seed: int = 2):
    """least_loaded snapshot, homogeneous nodes, pod_count ~ U{0..110}."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return homogeneous_nodes(n, rng, 110), uniform_pods(p, rng)


def config3(n: int = 5000, p: int = 1_000_000, seed: int = 3):
    """resource_balanced greedy, homogeneous nodes, pod_count ~ U{0..55}."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return homogeneous_nodes(n, rng, 55), uniform_pods(p, rng)


def config4(n: int = 50_000, p: int = 10_000_000, seed: int = 4):
    """resource_balanced snapshot, as config2 but larger."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return homogeneous_nodes(n, rng, 110), uniform_pods(p, rng)


def config5(n: int = 20_000, p: int = 2_000_000, seed: int = 5, infeasible: float = 0.30):
    """Mixed strategies, heterogeneous shapes, 2 % not Ready, 1 % max_pods == 0, and
    `infeasible` of the pods larger than any node (so they fit nowhere)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sh = rng.integers(0, len(SHAPES), n)
    cpu = np.array([s[0] for s in SHAPES], dtype=np.int64)[sh]
    mem = np.array([s[1] for s in SHAPES], dtype=np.int64)[sh]
    mp = np.array([s[2] for s in SHAPES], dtype=np.int64)[sh]
    mp[rng.random(n) < 0.01] = 0
    pc = (rng.random(n) * (mp + 1)).astype(np.int64)
    ready = (rng.random(n) >= 0.02).astype(np.uint8)
    u = usage_a3(pc, mp)
    nodes = NodeTable.from_columns(alloc_cpu_milli=cpu, alloc_mem_bytes=mem, pod_count=pc,
                                   max_pods=mp, ready=ready, cpu_usage_pct=u, mem_usage_pct=u.copy())
    pods = uniform_pods(p, rng)
    big = rng.random(p) < infeasible
    kind = rng.random(p) < 0.5
    pods.req_cpu_milli[big & kind] = rng.integers(100_000, 400_001, int((big & kind).sum()))
    pods.req_mem_bytes[big & ~kind] = rng.integers(800, 4097, int((big & ~kind).sum())) * GIB
    pods.strategy_ids = rng.integers(0, 3, p, dtype=np.int64)
    return nodes, pods


def all_equal_nodes(n: int, pod_count: int = 7) -> NodeTable:
    """Every node identical: every score ties, the expected winner is the first eligible node."""
    cpu, mem, mp = HOMOGENEOUS_SHAPE
    pc = np.full(n, pod_count, dtype=np.int64)
    mpa = np.full(n, mp, dtype=np.int64)
    u = usage_a3(pc, mpa)
    return NodeTable.from_columns(alloc_cpu_milli=np.full(n, cpu), alloc_mem_bytes=np.full(n, mem),
                                  pod_count=pc, max_pods=mpa, ready=np.ones(n, np.uint8),
                                  cpu_usage_pct=u, mem_usage_pct=u.copy())
