"""Read sharding across the GPUs of one box.

The path has no cross-read state (alnmain.c:330-479 keeps `best[]`, `m[]`, `cig[]` per read), so a batch is
split into `world` CONTIGUOUS read ranges, one per rank, every rank maps its range against its own replica of
the index, and the host concatenates the per-rank outputs in rank order -- the result is byte-identical to the
one-GPU run (alnmain.c:488 prints in read order).  There is no data-path collective.  The only collective is the
end-of-run sum of the per-shard counters behind the reference's `Sensitivity: valid/total` line (alnmain.c:541):
one all-reduce of a few int64 (NCCL over NVLink on the GPUs, gloo in the CPU tests).
"""
import numpy as np

STAT_KEYS = ("total", "valid", "n_flag4", "n_flag16", "sum_ed", "dp_cells", "seeds", "sa_hits", "bs_steps",
             "overflow_reads")


def shard_range(n, rank, world):
    """[begin, end) of rank's contiguous slice of n reads; the first n % world ranks get one extra read."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(int(n), world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def reduce_stats(stats, device=None, group=None):
    """Sum the as_stats counters of every rank (the one collective of the design).  No-op without a process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return dict(stats)
    keys = [k for k in STAT_KEYS if k in stats] + sorted(k for k in stats if k not in STAT_KEYS)
    t = torch.tensor([int(stats[k]) for k in keys], dtype=torch.int64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return dict(zip(keys, [int(x) for x in t.tolist()]))


def max_over_ranks(value, device=None, group=None):
    """Device time of a multi-GPU step = the slowest rank's."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def gather_in_rank_order(local, dst=0, group=None):
    """Concatenate per-rank numpy arrays (results, CIGAR columns, SAM bytes) on rank `dst`, rank 0's first.
    Host-side and off the timed path: this is the `cat` of the per-rank SAM streams."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    parts = [None] * world if rank == dst else None
    dist.gather_object(local, parts, dst=dst, group=group)
    if rank != dst:
        return None
    return np.concatenate(parts) if isinstance(local, np.ndarray) else b"".join(parts)


def bind_to_gpu_numa(device_index):
    """Pin this process (and so the pinned host buffers it allocates afterwards) to the NUMA node of its GPU.

    One process per GPU streams its read shard over that GPU's PCIe link; with eight processes the copies of
    the GPUs on the other socket otherwise cross the inter-socket link both ways.  Returns the node, or None when
    the box does not say (one node, no sysfs entry) -- then nothing is changed."""
    import os
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = parse_cpulist(f.read())
        allowed = os.sched_getaffinity(0)
        cpus = cpus & allowed if cpus & allowed else cpus
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError, RuntimeError):
        return None


def parse_cpulist(text):
    """'0-3,8,10-11' -> {0,1,2,3,8,10,11} (the kernel's cpulist format)."""
    out = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        out.update(range(int(lo), int(hi or lo) + 1))
    return out
