"""Print the GPU -> NUMA node map of the box and the node CPU lists (what shard.bind_host_to_device acts on)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from __graft_entry__ import load
load()
from gasma_b200 import shard
for d in range(torch.cuda.device_count()):
    print("gpu", d, "numa", shard.device_numa_node(d))
base = "/sys/devices/system/node"
for n in sorted(os.listdir(base)):
    if n.startswith("node"):
        print(n, open(os.path.join(base, n, "cpulist")).read().strip(), open(os.path.join(base, n, "meminfo")).readline().strip())
print("affinity", len(os.sched_getaffinity(0)))
