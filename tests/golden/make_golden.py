#!/usr/bin/env python3
"""Generate the golden vectors under tests/golden/ by RUNNING THE REFERENCE ITSELF.

This script is the only thing in the repository that touches /root/reference, and it
only runs in the build container (the GPU box has no /root/reference).  It imports the
unmodified reference ``scheduler.py`` with two shims:

  * ``HUGGINGFACE_TOKEN`` set to a dummy value (the module exits at import otherwise,
    scheduler.py:62-66);
  * stub ``kubernetes`` modules registered in ``sys.modules`` (scheduler.py:17-18).

and records, as JSON, what the reference computes for the fallback placement path:

  kat_scores.json      best_score bits of ``_fallback_decision`` (scheduler.py:521-559)
                       on single-node tables, captured with ``sys.settrace`` from the
                       function's own local ``best_score`` - no restatement involved.
  decisions.json       chosen index + best_score bits on seeded random clusters (ties,
                       not-Ready nodes, max_pods == 0, over-full nodes), 3 strategies
                       plus an unknown strategy string.
  greedy.json          bind sequences of the reference's own loop: real
                       ``ContextManager.get_node_metrics`` (scheduler.py:121-170) over a
                       fake CoreV1Api, real ``_fallback_decision``, fake bind.
  parsers.json         ``_parse_cpu`` / ``_parse_memory`` (scheduler.py:172-187) and
                       ``_convert_pod_to_spec`` (scheduler.py:731-764) outputs.
  ingest.json          a few hundred pods / node rows with randomly spelled quantity strings
                       (incl. spellings float() accepts or rejects in surprising ways) through
                       ``_convert_pod_to_spec`` and ``get_node_metrics`` (scheduler.py:121-170).
  config1_trace.json   BASELINE config 1: 3-node cluster x ai-test-pods.yaml through
                       ``CustomScheduler._schedule_pod`` with the LLM call failing.

Run:  python tests/golden/make_golden.py           (rewrites the JSON files in place)
"""
from __future__ import annotations

import asyncio
import json
import os
import struct
import sys
import types
from pathlib import Path

import numpy as np

REF_DIR = Path(os.environ.get("KP_REFERENCE_DIR", "/root/reference"))
OUT_DIR = Path(__file__).resolve().parent


# --------------------------------------------------------------------------------------
# reference import (unmodified file, two shims)
# --------------------------------------------------------------------------------------
def import_reference():
    os.environ.setdefault("HUGGINGFACE_TOKEN", "dummy-token-for-offline-import")
    os.environ.setdefault("LOG_LEVEL", "CRITICAL")

    class ApiException(Exception):
        def __init__(self, status=0, reason=""):
            super().__init__(f"({status}) {reason}")
            self.status, self.reason = status, reason

    class _Obj:
        def __init__(self, **kw):
            self.__dict__.update(kw)

    k = types.ModuleType("kubernetes")
    kc = types.ModuleType("kubernetes.client")
    kcfg = types.ModuleType("kubernetes.config")
    kw = types.ModuleType("kubernetes.watch")
    kr = types.ModuleType("kubernetes.client.rest")
    kr.ApiException = ApiException
    kcfg.load_kube_config = lambda *a, **kw_: None
    kc.CoreV1Api = lambda *a, **kw_: None
    kc.V1Binding = _Obj
    kc.V1ObjectMeta = _Obj
    kc.V1ObjectReference = _Obj
    kc.rest = kr
    kw.Watch = _Obj
    k.client, k.config, k.watch = kc, kcfg, kw
    sys.modules.update({
        "kubernetes": k, "kubernetes.client": kc, "kubernetes.config": kcfg,
        "kubernetes.watch": kw, "kubernetes.client.rest": kr,
    })
    sys.path.insert(0, str(REF_DIR))
    import scheduler  # noqa: E402  (the reference)
    return scheduler


def f64_bits(x):
    if x is None:          # empty node list: the function returns before best_score exists
        return None
    return struct.pack(">d", float(x)).hex()


# --------------------------------------------------------------------------------------
# helpers that call the reference function and capture its local best_score
# --------------------------------------------------------------------------------------
class ScoreTap:
    """sys.settrace hook: grabs best_node / best_score locals at _fallback_decision return."""

    def __init__(self, sched):
        self.code = sched.HuggingFaceClient._fallback_decision.__code__
        self.best_score = None
        self.best_node = None

    def _local(self, frame, event, arg):
        if event == "return":
            self.best_score = frame.f_locals.get("best_score")
            self.best_node = frame.f_locals.get("best_node")
        return self._local

    def _global(self, frame, event, arg):
        if frame.f_code is self.code:
            return self._local
        return None

    def decide(self, sched, nodes, strategy, reason="golden"):
        sched.CONFIG.setdefault("fallback", {})["strategy"] = strategy
        self.best_score, self.best_node = None, None
        sys.settrace(self._global)
        try:
            dec = sched.HuggingFaceClient._fallback_decision(None, nodes, reason)
        finally:
            sys.settrace(None)
        return dec


READY = [{"type": "Ready", "status": "True", "reason": "KubeletReady"}]
COND_VARIANTS = {
    "ready": READY,
    "false": [{"type": "Ready", "status": "False", "reason": "x"}],
    "unknown": [{"type": "Ready", "status": "Unknown", "reason": "x"}],
    "none": [],
    "other_only": [{"type": "MemoryPressure", "status": "True", "reason": "x"}],
    "ready_second": [{"type": "DiskPressure", "status": "False", "reason": "x"},
                     {"type": "Ready", "status": "True", "reason": "x"}],
}


def usage_a3(pc: int, mp: int):
    # the reference's own expression (scheduler.py:150-151), evaluated by CPython
    return (pc / mp) * 50 if mp > 0 else 0


def mk_node(sched, i, pc, mp, cu=None, mu=None, cond="ready", cpu=8.0, mem=15.616012573242188):
    u = usage_a3(pc, mp)
    return sched.NodeMetrics(
        name=f"node-{i}", cpu_usage_percent=u if cu is None else cu,
        memory_usage_percent=u if mu is None else mu, available_cpu=cpu,
        available_memory=mem, pod_count=pc, max_pods=mp, labels={}, taints=[],
        conditions=COND_VARIANTS[cond])


STRATS = ["resource_balanced", "least_loaded", "round_robin", "no_such_strategy"]


def gen_kat_scores(sched, tap):
    rows = []
    pairs = [(2, 110), (3, 110), (8, 110), (55, 110), (1, 3), (0, 0), (109, 110), (110, 110),
             (0, 110), (1, 110), (1, 2), (2, 6), (341, 110), (223, 110), (222, 110), (340, 110),
             (7, 30), (29, 58), (250, 250), (125, 250), (17, 58), (0, 1), (1, 1), (5, 0)]
    rng = np.random.Generator(np.random.PCG64(11))
    for _ in range(200):
        mp = int(rng.choice([30, 58, 110, 250, 1, 7, 1000]))
        pairs.append((int(rng.integers(0, mp + 1)), mp))
    for pc, mp in pairs:
        node = mk_node(sched, 0, pc, mp)
        row = {"pc": pc, "mp": mp, "usage_bits": f64_bits(node.cpu_usage_percent)}
        for s in STRATS[:3]:
            dec = tap.decide(sched, [node], s)
            row[s] = {"selected": dec.selected_node != "",
                      "score_bits": f64_bits(tap.best_score),
                      "score_is_int": isinstance(tap.best_score, int)}
        rows.append(row)
    # custom usage values that do not follow A3 (direct NodeMetrics input)
    custom = []
    for cu, mu, pc, mp in [(0.0, 0.0, 0, 110), (100.0, 100.0, 0, 110), (150.0, 20.0, 4, 110),
                           (201.0, 0.5, 4, 110), (33.3, 66.6, 10, 30), (99.999, 1e-9, 1, 58),
                           (250.0, 250.0, 500, 110), (12.5, 87.5, 0, 0), (400.0, 0.0, 0, 10)]:
        node = mk_node(sched, 0, pc, mp, cu=cu, mu=mu)
        row = {"pc": pc, "mp": mp, "cu": cu, "mu": mu}
        for s in STRATS[:3]:
            dec = tap.decide(sched, [node], s)
            row[s] = {"selected": dec.selected_node != "",
                      "score_bits": f64_bits(tap.best_score)}
        custom.append(row)
    return {"a3_usage": rows, "custom_usage": custom}


def gen_decisions(sched, tap):
    cases = []
    rng = np.random.Generator(np.random.PCG64(7))
    shapes = [30, 58, 110, 250]

    def record(tag, spec):
        nodes = [mk_node(sched, i, *sp) for i, sp in enumerate(spec)]
        out = {}
        for s in STRATS:
            dec = tap.decide(sched, nodes, s, reason=f"r-{tag}")
            idx = -1 if tap.best_node is None else nodes.index(tap.best_node)
            assert (dec.selected_node == "") == (idx == -1)
            if idx >= 0:
                assert dec.selected_node == nodes[idx].name
            out[s] = {"idx": idx, "score_bits": f64_bits(tap.best_score),
                      "confidence": dec.confidence, "reasoning": dec.reasoning,
                      "fallback_needed": dec.fallback_needed}
        rows = []
        for sp, n in zip(spec, nodes):
            row = {"pc": sp[0], "mp": sp[1], "cu_bits": f64_bits(n.cpu_usage_percent),
                   "mu_bits": f64_bits(n.memory_usage_percent), "cond": sp[4] if len(sp) > 4 else "ready"}
            if len(sp) > 5:                     # allocatable fields the decision never reads (scheduler.py:531-549)
                row["avail_cpu_bits"] = f64_bits(n.available_cpu)
                row["avail_mem_bits"] = f64_bits(n.available_memory)
            rows.append(row)
        cases.append({"tag": tag, "nodes": rows, "expect": out})

    # behavioural known answers from SURVEY.md §8(c)
    record("kat-10-5-5", [(10, 110), (5, 110), (5, 110)])
    record("kat-half-zero", [(55, 110), (0, 0)])
    record("kat-r04-zero", [(44, 110), (0, 0)])
    record("kat-341", [(341, 110)])
    record("kat-223", [(223, 110)])
    record("kat-all-notready", [(1, 110, None, None, "false"), (2, 110, None, None, "unknown"),
                                (3, 110, None, None, "none"), (0, 110, None, None, "other_only")])
    record("kat-ready-second", [(9, 110, None, None, "false"), (9, 110, None, None, "ready_second"),
                                (9, 110)])
    record("kat-exact-ties", [(1, 2), (55, 110), (2, 4), (1, 3), (2, 6)])
    record("kat-empty", [])
    # seeded random clusters
    for c in range(60):
        n = int(rng.integers(1, 48))
        spec = []
        homogeneous = c % 3 == 0
        for _ in range(n):
            mp = 110 if homogeneous else int(rng.choice(shapes))
            r = rng.random()
            if r < 0.04:
                mp = 0
            pc = int(rng.integers(0, (mp if mp > 0 else 5) + 1))
            if rng.random() < 0.03:
                pc = int(mp * rng.uniform(2.0, 3.5)) if mp > 0 else 3
            cond = "ready"
            r = rng.random()
            if r < 0.08:
                cond = ["false", "unknown", "none", "other_only"][int(rng.integers(0, 4))]
            elif r < 0.12:
                cond = "ready_second"
            cu = mu = None
            if c % 5 == 4 and rng.random() < 0.5:       # usage that does not follow A3
                cu = float(np.round(rng.uniform(0, 120), 3))
                mu = float(np.round(rng.uniform(0, 120), 3))
            spec.append((pc, mp, cu, mu, cond))
        record(f"rand-{c}", spec)
    # inputs the reference accepts because it never reads them (available_cpu / available_memory) or
    # because plain float arithmetic copes (inf / nan / huge / negative usage)
    inf, nan = float("inf"), float("nan")
    record("odd-alloc", [(10, 110, None, None, "ready", inf, nan), (5, 110, None, None, "ready", 1e300, -4.0),
                         (5, 110, None, None, "ready", nan, inf), (7, 110, None, None, "ready", -inf, 2.0 ** 60)])
    record("odd-alloc-notready", [(1, 110, None, None, "false", nan, nan), (3, 110, None, None, "ready", -1e300, 1e300)])
    record("odd-usage", [(10, 110, inf, 3.0), (5, 110, nan, nan), (6, 110, -5.0, -7.5), (4, 110, 1e300, 0.0),
                         (9, 110, 20.0, 20.0), (8, 110, -inf, 1.0)])
    record("odd-usage-only", [(10, 110, nan, 3.0), (5, 110, 3.0, nan), (2, 110, inf, inf)])
    return cases


# --------------------------------------------------------------------------------------
# fake K8s API server: the complete surface scheduler.py touches
# --------------------------------------------------------------------------------------
class _NS:
    def __init__(self, **kw):
        self.__dict__.update(kw)


class FakeCoreV1:
    def __init__(self, node_specs):
        # node_specs: list of dict(name, cpu, memory, pods, n_running, cond)
        self.node_specs = node_specs
        self.pods_on = {ns["name"]: ns["n_running"] for ns in node_specs}
        self.binds = []

    def list_node(self):
        items = []
        for ns in self.node_specs:
            conds = [_NS(type=c["type"], status=c["status"], reason=c.get("reason"))
                     for c in COND_VARIANTS[ns.get("cond", "ready")]]
            alloc = {"cpu": ns["cpu"], "memory": ns["memory"], "pods": ns["pods"]}
            items.append(_NS(metadata=_NS(name=ns["name"], labels=None),
                             spec=_NS(taints=None),
                             status=_NS(conditions=conds, capacity=dict(alloc), allocatable=alloc)))
        return _NS(items=items)

    def list_pod_for_all_namespaces(self, field_selector=None, **kw):
        name = field_selector.split("=", 1)[1]
        return _NS(items=[None] * self.pods_on[name])

    def create_namespaced_binding(self, namespace, body, _preload_content=False):
        target = body.target.name
        self.pods_on[target] += 1
        self.binds.append(target)


def gen_greedy(sched, tap):
    cases = []
    rng = np.random.Generator(np.random.PCG64(13))

    def run(tag, node_specs, n_pods, strategy):
        api = FakeCoreV1([dict(ns) for ns in node_specs])
        cm = sched.ContextManager.__new__(sched.ContextManager)
        cm.v1 = api
        seq, scores = [], []
        first_table = None
        for _ in range(n_pods):
            table = cm.get_node_metrics()              # real reference code, A3 usage formula
            if first_table is None:
                first_table = table
            dec = tap.decide(sched, table, strategy)
            names = [t.name for t in table]
            idx = names.index(dec.selected_node) if dec.selected_node else -1
            seq.append(idx)
            scores.append(f64_bits(tap.best_score))
            if idx >= 0:                               # scheduler.py:716-721 -> bind
                api.pods_on[names[idx]] += 1
        final = [api.pods_on[ns["name"]] for ns in node_specs]
        cases.append({
            "tag": tag, "strategy": strategy, "n_pods": n_pods,
            "nodes": node_specs,
            "parsed": [{"available_cpu_bits": f64_bits(t.available_cpu),
                        "available_memory_bits": f64_bits(t.available_memory),
                        "max_pods": t.max_pods, "pod_count": t.pod_count,
                        "cu_bits": f64_bits(t.cpu_usage_percent)} for t in first_table],
            "idx_seq": seq, "score_bits_seq": scores, "final_pod_count": final})

    cfg1 = [dict(name="minikube", cpu="8", memory="16374576Ki", pods="110", n_running=8),
            dict(name="minikube-m02", cpu="8", memory="16374576Ki", pods="110", n_running=2),
            dict(name="minikube-m03", cpu="8", memory="16374576Ki", pods="110", n_running=2)]
    for s in STRATS[:3]:
        run(f"config1-{s}", cfg1, 12, s)
    shapes = [("4", "16Gi", "30"), ("16", "64Gi", "58"), ("32", "128Gi", "110"),
              ("96", "768Gi", "250"), ("7910m", "32780300Ki", "110"), ("2", "4096Mi", "8")]
    for c in range(12):
        n = int(rng.integers(2, 24))
        specs = []
        for i in range(n):
            cpu, mem, pods = shapes[int(rng.integers(0, len(shapes)))] if c % 2 else shapes[2]
            if rng.random() < 0.05:
                pods = "0"
            cond = "ready" if rng.random() > 0.1 else "false"
            mp = int(pods)
            specs.append(dict(name=f"n{i}", cpu=cpu, memory=mem, pods=pods,
                              n_running=int(rng.integers(0, max(mp, 1) + 1)), cond=cond))
        for s in STRATS[:3]:
            run(f"rand-{c}-{s}", specs, int(rng.integers(20, 120)), s)
    # a tiny cluster driven past max_pods: the reference never stops at pc == mp, and a
    # node stops winning only once its score falls to <= -1 (scheduler.py:529,547)
    tiny = [dict(name="a", cpu="2", memory="4Gi", pods="3", n_running=0),
            dict(name="b", cpu="2", memory="4Gi", pods="2", n_running=1)]
    for s in STRATS[:3]:
        run(f"overfill-{s}", tiny, 24, s)
    return cases


def gen_parsers(sched):
    cm = sched.ContextManager.__new__(sched.ContextManager)
    cpu_in = ["8", "250m", "500m", "100m", "7910m", "0", "0.5", "1.5", "32", "96", "4000m",
              "1m", "123456m", "2.25", "16"]
    mem_in = ["16374576Ki", "256Mi", "512Mi", "128Mi", "16Gi", "768Gi", "0", "1073741824",
              "32780300Ki", "1.5Gi", "100Mi", "4096Mi", "17179869184", "1Ki", "0.5Mi"]
    bad_mem = ["1G", "1M", "1K", "1Ti", "1e3Mi"[:0] or "12x"]
    out = {"parse_cpu": [{"in": s, "bits": f64_bits(cm._parse_cpu(s))} for s in cpu_in],
           "parse_memory": [{"in": s, "bits": f64_bits(cm._parse_memory(s))} for s in mem_in],
           "parse_memory_errors": []}
    for s in bad_mem:
        try:
            cm._parse_memory(s)
            out["parse_memory_errors"].append({"in": s, "error": None})
        except Exception as e:  # noqa: BLE001
            out["parse_memory_errors"].append({"in": s, "error": type(e).__name__})

    cs = sched.CustomScheduler.__new__(sched.CustomScheduler)
    pods = [
        [("250m", "256Mi")], [("500m", "512Mi")], [("100m", "128Mi")],
        [("100m", "64Mi"), ("200m", "64Mi")],                       # 0.1 + 0.2 in fp64
        [("1", "1Gi"), ("0.5", "512Mi"), ("250m", "1048576Ki")],
        [("100m", "1073741824")],                                   # plain bytes -> ignored (A4 quirk)
        [(None, None)], [("", "")], [],
        [("300m", "100Mi"), ("300m", "100Mi"), ("300m", "100Mi")],
        [(2, 1024)],                                                # non-string values
    ]
    conv = []
    for i, containers in enumerate(pods):
        cl = []
        for cpu, mem in containers:
            req = {}
            if cpu is not None:
                req["cpu"] = cpu
            if mem is not None:
                req["memory"] = mem
            cl.append(_NS(resources=_NS(requests=req) if req or cpu is None else None))
        pod = _NS(metadata=_NS(name=f"p{i}", namespace="default"),
                  spec=_NS(containers=cl, node_selector=None, tolerations=None, priority=None))
        ps = cs._convert_pod_to_spec(pod)
        conv.append({"containers": [[c, m] for c, m in containers],
                     "cpu_bits": f64_bits(ps.cpu_request), "mem_bits": f64_bits(ps.memory_request),
                     "priority": ps.priority})
    out["convert_pod_to_spec"] = conv
    return out


# --------------------------------------------------------------------------------------
# ingest vectors: quantity strings through the reference's own parsers, in bulk, including
# every spelling class the device grammar accepts, defers to the host, or must reject
# --------------------------------------------------------------------------------------
ODD_NUMBERS = ["", "+5", "-5", "-0", "1e3", "1E-2", " 5", "5 ", "1_0", "\u0663", "inf", "nan", "Infinity",
               "0.1234567890123456789", "12345678901234567890", "9007199254740993", ".5", "5.", ".",
               "1.5.2", "0x10", "--1", "1,5", "00012", "0.000", "3.14159", "007.500", "1e", "e5"]


def _rand_number(rng):
    k = int(rng.integers(0, 10))
    if k == 0 and rng.random() < 0.4:
        return ODD_NUMBERS[int(rng.integers(0, len(ODD_NUMBERS)))]
    if k <= 4:
        return str(int(rng.integers(0, 200000)))
    if k <= 6:
        return f"{int(rng.integers(0, 64))}.{int(rng.integers(0, 1000)):0{int(rng.integers(1, 4))}d}"
    if k == 7:
        return f"{rng.random() * 100:.{int(rng.integers(4, 12))}f}"
    if k == 8:
        return str(int(rng.integers(0, 2 ** 40)))
    return f"0.{int(rng.integers(0, 10))}"


def gen_ingest(sched):
    rng = np.random.Generator(np.random.PCG64(29))
    cs = sched.CustomScheduler.__new__(sched.CustomScheduler)
    def sfx(good, bad):
        return bad[int(rng.integers(0, len(bad)))] if rng.random() < 0.05 else good[int(rng.integers(0, len(good)))]

    cpu_good, cpu_bad = ["m", "m", "m", "", ""], ["M", "mm", " m"]
    mem_good, mem_bad = ["Ki", "Mi", "Gi", "Mi", "Gi", "", ""], ["G", "K", "Ti", "ki", "i", "Gi "]
    pods = []
    for i in range(420):
        containers = []
        for _ in range(int(rng.integers(0, 5)) if i % 7 else 1):
            r = rng.random()
            if r < 0.06:
                containers.append(None)                      # container without resources.requests
                continue
            req = {}
            if rng.random() > 0.05:
                req["cpu"] = _rand_number(rng) + sfx(cpu_good, cpu_bad)
            if rng.random() > 0.05:
                req["memory"] = _rand_number(rng) + sfx(mem_good, mem_bad)
            if rng.random() < 0.02:
                req["cpu"] = int(rng.integers(0, 8))         # YAML integer instead of a string
            containers.append(req)
        cl = [_NS(resources=_NS(requests=req) if req is not None else None) for req in containers]
        pod = _NS(metadata=_NS(name=f"p{i}", namespace="default"),
                  spec=_NS(containers=cl, node_selector=None, tolerations=None, priority=None))
        rec = {"containers": containers}
        try:
            ps = cs._convert_pod_to_spec(pod)
            rec.update(cpu_bits=f64_bits(ps.cpu_request), mem_bits=f64_bits(ps.memory_request), error=None)
        except Exception as e:  # noqa: BLE001
            rec.update(cpu_bits=None, mem_bits=None, error=type(e).__name__)
        pods.append(rec)

    nodes = []
    for i in range(260):
        cpu = _rand_number(rng) + sfx(cpu_good, cpu_bad)
        mem = _rand_number(rng) + sfx(mem_good, mem_bad)
        pods_s = str(int(rng.integers(0, 300))) if rng.random() > 0.08 else ODD_NUMBERS[int(rng.integers(0, len(ODD_NUMBERS)))]
        spec = dict(name=f"n{i}", cpu=cpu, memory=mem, pods=pods_s, n_running=int(rng.integers(0, 260)))
        cm = sched.ContextManager.__new__(sched.ContextManager)
        cm.v1 = FakeCoreV1([dict(spec)])
        table = cm.get_node_metrics()                        # [] on any exception (scheduler.py:168-170)
        rec = dict(spec)
        if table:
            t = table[0]
            rec.update(ok=True, available_cpu_bits=f64_bits(t.available_cpu),
                       available_memory_bits=f64_bits(t.available_memory), max_pods=t.max_pods,
                       pod_count=t.pod_count, cu_bits=f64_bits(t.cpu_usage_percent),
                       mu_bits=f64_bits(t.memory_usage_percent))
        else:
            rec.update(ok=False)
        nodes.append(rec)
    return {"pods": pods, "nodes": nodes}


def gen_prompts(sched):
    """The computed node block of PromptEngine.construct_scheduling_prompt (scheduler.py:226-243)."""
    pe = sched.PromptEngine()
    rng = np.random.Generator(np.random.PCG64(41))
    out = []
    for c in range(4):
        nodes = []
        for i in range(int(rng.integers(1, 9))):
            mp = int(rng.choice([30, 58, 110, 250]))
            pc = int(rng.integers(0, mp + 1))
            u = usage_a3(pc, mp)
            cpu = float(rng.choice([4.0, 7.91, 16.0, 32.0, 96.0]))
            mem = float(rng.choice([16374576 / 1024 / 1024, 64.0, 128.0, 32780300 / 1024 / 1024]))
            nodes.append(dict(name=f"node-{c}-{i}", pc=pc, mp=mp, cu_bits=f64_bits(u), cpu=cpu, mem=mem))
        recs = [mk_node(sched, n["name"], n["pc"], n["mp"], cpu=n["cpu"], mem=n["mem"]) for n in nodes]
        for r, n in zip(recs, nodes):
            r.name = n["name"]
        pod = sched.PodSpec(name="p", namespace="default", cpu_request=0.25, memory_request=0.25,
                            node_selector={}, tolerations=[], affinity_rules={}, priority=0)
        prompt = pe.construct_scheduling_prompt(pod, recs)
        a = prompt.index("AVAILABLE NODES:")
        b = prompt.index("VALID NODE NAMES:")
        b = prompt.index("\n", b) + 1
        out.append({"nodes": nodes, "cluster_info": prompt[a:b]})
    return out


def gen_cache_keys(sched):
    """RequestCache._generate_key (scheduler.py:265-271) on a few clusters and pods."""
    cache = sched.RequestCache()
    rng = np.random.Generator(np.random.PCG64(43))
    out = []
    for c, n_nodes in enumerate([0, 1, 3, 40, 700]):
        nodes = []
        for i in rng.permutation(n_nodes):
            mp = int(rng.choice([30, 58, 110, 250]))
            pc = int(rng.integers(0, mp + 1))
            nodes.append(dict(name=f"ip-10-0-{int(i) // 250}-{int(i) % 250}.ec2.internal" if c % 2 else f"node-{int(i)}",
                              pc=pc, mp=mp, cu_bits=f64_bits(usage_a3(pc, mp))))
        recs = [mk_node(sched, 0, n["pc"], n["mp"]) for n in nodes]
        for r, n in zip(recs, nodes):
            r.name = n["name"]
        pods, keys = [], []
        for j in range(12):
            cpu = [0.25, 0.5, 0.1, 0.30000000000000004, 4.0, 1e-05, 123456.789, 0.0][j % 8]
            mem = [0.25, 0.5, 0.125, 1.0, 16.0, 0.0009765625, 1e22, 0.0][(j * 3) % 8]
            prio = [0, 1000, -5, 2000000000][j % 4]
            pod = sched.PodSpec(name=f"p{j}", namespace="default", cpu_request=cpu, memory_request=mem,
                                node_selector={}, tolerations=[], affinity_rules={}, priority=prio)
            pods.append(dict(cpu_bits=f64_bits(cpu), mem_bits=f64_bits(mem), priority=prio))
            keys.append(cache._generate_key(pod, recs))
        out.append({"nodes": nodes, "pods": pods, "keys": keys})
    return out


def gen_config1_trace(sched):
    """BASELINE config 1 through the reference's own _schedule_pod."""
    import yaml
    docs = list(yaml.safe_load_all((REF_DIR / "ai-test-pods.yaml").read_text()))
    api = FakeCoreV1([
        dict(name="minikube", cpu="8", memory="16374576Ki", pods="110", n_running=8),
        dict(name="minikube-m02", cpu="8", memory="16374576Ki", pods="110", n_running=2),
        dict(name="minikube-m03", cpu="8", memory="16374576Ki", pods="110", n_running=2)])
    sleeps = []
    real_sleep = sched.time.sleep
    sched.time.sleep = lambda s: sleeps.append(s)
    sched.CONFIG.setdefault("fallback", {})["strategy"] = "resource_balanced"
    try:
        cm = sched.ContextManager.__new__(sched.ContextManager)
        cm.v1 = api
        llm = sched.HuggingFaceClient(sched.LLM_MODEL, "dummy", sched.LLM_ENDPOINT)

        def boom(*a, **kw):
            raise RuntimeError("offline: no HuggingFace endpoint")
        llm.client.chat_completion = boom
        il = sched.IntegrationLayer.__new__(sched.IntegrationLayer)
        il.v1 = api
        cs = sched.CustomScheduler.__new__(sched.CustomScheduler)
        cs.context_manager, cs.prompt_engine = cm, sched.PromptEngine()
        cs.llm_client, cs.integration_layer = llm, il
        cs.running = True
        cs.stats = {"total_scheduled": 0, "llm_decisions": 0, "fallback_decisions": 0,
                    "failed_bindings": 0}
        specs = []
        for d in docs:
            c = d["spec"]["containers"][0]
            pod = _NS(metadata=_NS(name=d["metadata"]["name"], namespace=d["metadata"]["namespace"]),
                      spec=_NS(containers=[_NS(resources=_NS(requests=dict(c["resources"]["requests"])))],
                               node_selector=None, tolerations=None, priority=None))
            ps = cs._convert_pod_to_spec(pod)
            specs.append({"name": ps.name, "cpu_bits": f64_bits(ps.cpu_request),
                          "mem_bits": f64_bits(ps.memory_request),
                          "requests": dict(c["resources"]["requests"])})
            asyncio.run(cs._schedule_pod(pod))
    finally:
        sched.time.sleep = real_sleep
    table = cm.get_node_metrics()
    return {"pods": specs, "binds": api.binds, "sleeps": sleeps, "stats": cs.stats,
            "llm_stats": llm.get_stats(),
            "final_table": [{"name": t.name, "pod_count": t.pod_count,
                             "available_cpu_bits": f64_bits(t.available_cpu),
                             "available_memory_bits": f64_bits(t.available_memory),
                             "cu_bits": f64_bits(t.cpu_usage_percent)} for t in table]}


def main():
    sched = import_reference()
    tap = ScoreTap(sched)
    outputs = {
        "kat_scores.json": gen_kat_scores(sched, tap),
        "decisions.json": gen_decisions(sched, tap),
        "greedy.json": gen_greedy(sched, tap),
        "parsers.json": gen_parsers(sched),
        "config1_trace.json": gen_config1_trace(sched),
        "ingest.json": gen_ingest(sched),
        "prompts.json": gen_prompts(sched),
        "cache_keys.json": gen_cache_keys(sched),
    }
    for name, obj in outputs.items():
        (OUT_DIR / name).write_text(json.dumps(obj, indent=1, sort_keys=True) + "\n")
        print(f"wrote {name}: {(OUT_DIR / name).stat().st_size} bytes")


if __name__ == "__main__":
    main()
