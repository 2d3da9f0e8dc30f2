"""A few one-pod decisions through kp_session_decide, for profiling (GPU box): python tools/decide_one.py [nodes]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from k8s_llm_scheduler_b200 import placement, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
nodes, pods = synth.config5(n, 8, seed=3)
with placement.Session(0, max_nodes=max(n, 16), max_pods=16) as s:
    for j in range(8):
        print(s.decide(nodes, j % 3, 1, int(pods.req_cpu_milli[j]), int(pods.req_mem_bytes[j])))
