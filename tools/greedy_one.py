import sys
sys.path.insert(0, '/root/repo')
import torch
from k8s_llm_scheduler_b200 import placement, synth
nodes, pods = synth.config3(5000, 100000)
dn = placement.nodes_to_device(nodes, "cuda:0")
rc = torch.as_tensor(pods.req_cpu_milli).cuda(); rm = torch.as_tensor(pods.req_mem_bytes).cuda()
placement.greedy_place(dn, rc, rm, 0, 1)
torch.cuda.synchronize()
print("ok")
