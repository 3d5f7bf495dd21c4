"""Sharding of a run across the GPUs of one box (one process per GPU) and the path's only collective.

Pairs are independent (the reference's benchmark::run is a plain loop, GASMA/benchmark/benchmark_utils.h:373-385), so a
run of `total` pairs is cut into batches of `batch` pairs that are dealt round-robin to the ranks; every batch is
identified by the global index of its first pair, which is also what the counter-RNG generator keys on, so the union of
all ranks' work is the same set of pairs whatever the world size.  The only cross-rank quantity is the counter vector
(total, nw_correct, greedy_correct, greedy_coverage; benchmark_utils.h:238,253-258): one all-reduce(SUM) at the end
of the run -- NCCL over NVLink on the GPU box, gloo in the CPU tests.  No torch types cross the C ABI; this module is
process plumbing only."""
import os


def world():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when launched directly."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0")))


def batch_schedule(total, batch, rank, world_size):
    """[(first_pair, n_pairs), ...] of the batches this rank processes (round-robin by batch index)."""
    if total < 0 or batch <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad schedule arguments")
    n_batches = (total + batch - 1) // batch
    out = []
    for b in range(rank, n_batches, world_size):
        first = b * batch
        out.append((first, min(batch, total - first)))
    return out


def shard_range(total, rank, world_size):
    """Contiguous [lo, hi) slice of `total` pairs for this rank (host-fed batches: GPU g takes pairs [g*N/G, (g+1)*N/G))."""
    if not (0 <= rank < world_size):
        raise ValueError("bad rank")
    return (total * rank) // world_size, (total * (rank + 1)) // world_size


def reduce_counters(counters, group=None):
    """In-place SUM of an integer counter tensor over all ranks; a no-op without an initialised process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(counters, op=dist.ReduceOp.SUM, group=group)
    return counters


def max_over_ranks(value, device=None, group=None):
    """Max of a python float over all ranks (multi-GPU timings are reported as the slowest rank)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def device_numa_node(device_index, sysfs="/sys"):
    """NUMA node the GPU's PCIe root hangs off (None when the platform does not say)."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open(os.path.join(sysfs, "bus/pci/devices", bus, "numa_node")) as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def bind_host_to_device(device_index, sysfs="/sys"):
    """Pin the calling process to the CPUs of the GPU's NUMA node, so that the pinned host buffers it allocates next
    (first-touch placement) and the thread that feeds the host-pointer C-ABI calls sit next to that GPU's PCIe root.
    With one process per GPU and every process streaming ~300 MB per batch, buffers on the wrong socket make all the
    links share one socket's memory controllers and the inter-socket fabric.  Returns (node, n_cpus) or None when the
    topology is unknown or single-node (nothing is changed then)."""
    node = device_numa_node(device_index, sysfs)
    if node is None:
        return None
    try:
        nodes = [d for d in os.listdir(os.path.join(sysfs, "devices/system/node")) if d.startswith("node") and d[4:].isdigit()]
        if len(nodes) < 2:
            return None
        with open(os.path.join(sysfs, "devices/system/node/node%d/cpulist" % node)) as f:
            cpus = _parse_cpulist(f.read())
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node, len(cpus)
    except Exception:
        return None
