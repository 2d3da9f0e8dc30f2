"""BatchScheduler.schedule_batch == the reference's one-event-at-a-time loop
(_watch_pods filter scheduler.py:673-675, _schedule_pod :690-729), including binds that fail.
The sequential side is the oracle's literal loop; on CPU the batch uses the oracle's greedy
as its placement function (host logic only), the gpu-marked test uses the CUDA path."""
import copy
from types import SimpleNamespace as NS

import numpy as np
import pytest

from k8s_llm_scheduler_b200.batch import BatchScheduler, container_requests
from k8s_llm_scheduler_b200.tables import PodTable
from oracle import c_oracle, py_oracle

READY = [{"type": "Ready", "status": "True"}]
NOT_READY = [{"type": "Ready", "status": "False"}]


def make_cluster(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    shapes = [("4", "16Gi", "30"), ("16", "64Gi", "58"), ("32", "128Gi", "110")]
    nodes = []
    for i in range(n):
        cpu, mem, mp = shapes[int(rng.integers(0, 3))]
        nodes.append(py_oracle.node_row(f"n{i}", cpu, mem, mp, int(rng.integers(0, int(mp))),
                                        READY if rng.random() > 0.1 else NOT_READY))
    return nodes


def make_pods(p, seed, scheduler="ai-llama-scheduler"):
    rng = np.random.Generator(np.random.PCG64(seed))
    pods = []
    for j in range(p):
        cs = [NS(resources=NS(requests={"cpu": f"{int(rng.integers(1, 40)) * 50}m",
                                        "memory": f"{int(rng.integers(1, 64)) * 64}Mi"}))
              for _ in range(int(rng.integers(1, 3)))]
        if j % 11 == 0:
            cs.append(NS(resources=None))
        pods.append(NS(metadata=NS(name=f"p{j}", namespace="default"),
                       spec=NS(containers=cs, scheduler_name=scheduler, node_name=None),
                       status=NS(phase="Pending")))
    return pods


def sequential_reference(nodes, pods, strategy, filter_mode, fail):
    """One pod at a time: decide on the current table, bind, commit only when the bind succeeded."""
    nodes = copy.deepcopy(nodes)
    stats = {"total_scheduled": 0, "llm_decisions": 0, "fallback_decisions": 0, "failed_bindings": 0}
    names, binds = [], []
    for j, pod in enumerate(pods):
        c, m = py_oracle.pod_requests(container_requests(pod))
        spec = py_oracle.PodSpec(pod.metadata.name, "default", c, m)
        dec = py_oracle.fallback_decision(nodes, "r", strategy, spec, filter_mode)
        stats["fallback_decisions"] += 1
        names.append(dec.selected_node)
        if not dec.selected_node:
            stats["failed_bindings"] += 1
            continue
        if j in fail:
            stats["failed_bindings"] += 1
            continue
        stats["total_scheduled"] += 1
        binds.append((pod.metadata.name, dec.selected_node))
        py_oracle.commit(next(nd for nd in nodes if nd.name == dec.selected_node))
    return names, binds, stats


def oracle_greedy_fn(nodes, pods, sid, filter_mode):
    return c_oracle.greedy(nodes.columns(), pods.req_cpu_milli, pods.req_mem_bytes, strategy=sid,
                           filter_mode=filter_mode)[0]


def host_pod_table(requests):
    return PodTable.from_container_requests(requests, exact=False)


def run_batch(sched, nodes, pods, fail):
    binds = []

    def bind(name, namespace, node):
        assert namespace == "default"
        j = int(name[1:])
        if j in fail:
            return False
        binds.append((name, node))
        return True

    out = sched.schedule_batch(pods, nodes, bind, reason="r")
    return [d.selected_node for _, d, _ in out], binds, out


@pytest.mark.parametrize("strategy", ["resource_balanced", "least_loaded", "round_robin", "something_else"])
@pytest.mark.parametrize("filter_mode", [0, 1])
@pytest.mark.parametrize("fail", [set(), {0}, {3, 4, 17, 58}, set(range(0, 60, 2))])
def test_batch_equals_sequential_loop_host_logic(strategy, filter_mode, fail):
    nodes, pods = make_cluster(12, 5), make_pods(60, 6)
    sched = BatchScheduler("ai-llama-scheduler", {"fallback": {"strategy": strategy}}, filter_mode=filter_mode,
                           greedy_fn=oracle_greedy_fn, pod_table_fn=host_pod_table)
    got_names, got_binds, out = run_batch(sched, nodes, pods, fail)
    want_names, want_binds, want_stats = sequential_reference(nodes, pods, strategy, filter_mode, fail)
    assert got_names == want_names
    assert got_binds == want_binds
    assert sched.stats == want_stats
    for (pod, dec, ok), name in zip(out, want_names):
        assert dec.fallback_needed is True
        assert (dec.confidence, dec.reasoning) == ((0.4, f"Fallback ({strategy}): r") if name else (0.0, "Fallback failed: r"))
        assert ok == (bool(name) and int(pod.metadata.name[1:]) not in fail)


def test_event_filter_and_empty_inputs():
    sched = BatchScheduler("ai-llama-scheduler", greedy_fn=oracle_greedy_fn, pod_table_fn=host_pod_table)
    pod = make_pods(1, 1)[0]
    assert sched.needs_scheduling(pod)
    pod.spec.node_name = "n1"
    assert not sched.needs_scheduling(pod)
    other = make_pods(1, 1, scheduler="default-scheduler")[0]
    assert not sched.needs_scheduling(other)
    running = make_pods(1, 1)[0]
    running.status.phase = "Running"
    assert not sched.needs_scheduling(running)
    assert sched.schedule_batch([], make_cluster(3, 1), lambda *a: True) == []
    assert sched.schedule_batch(make_pods(2, 1), [], lambda *a: True) == []          # scheduler.py:694-696
    assert sched.stats["fallback_decisions"] == 0
    # all nodes not ready: every pod gets the in-band "no node" record and counts as a failed binding
    dead = [py_oracle.node_row(f"n{i}", "4", "16Gi", "30", 1, NOT_READY) for i in range(3)]
    out = sched.schedule_batch(make_pods(4, 2), dead, lambda *a: True, reason="x")
    assert [d.selected_node for _, d, _ in out] == [""] * 4
    assert sched.stats == {"total_scheduled": 0, "llm_decisions": 0, "fallback_decisions": 4, "failed_bindings": 4}


def test_malformed_quantity_raises_like_convert_pod_to_spec():
    sched = BatchScheduler("s", greedy_fn=oracle_greedy_fn, pod_table_fn=host_pod_table)
    pods = make_pods(3, 3)
    pods[1].spec.containers[0].resources.requests["memory"] = "12xMi"
    with pytest.raises(ValueError):
        sched.schedule_batch(pods, make_cluster(3, 1), lambda *a: True)


@pytest.mark.gpu
@pytest.mark.parametrize("filter_mode", [0, 1])
def test_batch_equals_sequential_loop_on_the_gpu(filter_mode):
    nodes, pods = make_cluster(40, 8), make_pods(400, 9)
    fail = {5, 6, 200, 399}
    for strategy in ("resource_balanced", "least_loaded", "round_robin"):
        sched = BatchScheduler("ai-llama-scheduler", {"fallback": {"strategy": strategy}}, filter_mode=filter_mode,
                               max_nodes=64, max_pods=512)
        try:
            got_names, got_binds, _ = run_batch(sched, nodes, pods, fail)
        finally:
            sched.close()
        want_names, want_binds, want_stats = sequential_reference(nodes, pods, strategy, filter_mode, fail)
        assert got_names == want_names
        assert got_binds == want_binds
        assert sched.stats == want_stats
