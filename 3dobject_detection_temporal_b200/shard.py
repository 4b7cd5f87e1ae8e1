"""Frame sharding across the GPUs of one box (one process per GPU).

Frames are independent -- no state crosses frames -- so the path has no data-path
collective; each rank voxelizes its own frames exactly like the reference's
``DistributedSampler`` hands each rank its own frame indices
(second/pytorch/train_new_re.py:325-346)."""


def shard_frames(num_frames, rank, world_size, contiguous=False):
    """Frame indices owned by ``rank``.  Default is the DistributedSampler
    interleave (frame i -> rank i mod world_size)."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    if contiguous:
        per = (num_frames + world_size - 1) // world_size
        return list(range(min(rank * per, num_frames), min((rank + 1) * per, num_frames)))
    return list(range(rank, num_frames, world_size))


def gather_counts(local_points, local_voxels, group=None):
    """Sum of (points, voxels) over ranks -- bookkeeping outside the timed region."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return int(local_points), int(local_voxels)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor([int(local_points), int(local_voxels)], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return int(t[0].item()), int(t[1].item())


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def device_numa_node(device_index):
    """NUMA node the GPU hangs off (sysfs), or None when the box does not say."""
    import torch
    p = torch.cuda.get_device_properties(device_index)
    try:
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as fh:
            node = int(fh.read().strip())
    except (AttributeError, OSError, ValueError):
        return None
    return node if node >= 0 else None


def bind_host_to_device(device_index):
    """Run this process on the CPUs of the GPU's NUMA node, so the pinned host buffers it
    allocates afterwards (first touch) and its copies stay on the socket the GPU's PCIe
    root belongs to.  With one process per GPU the host legs of different ranks then stop
    crossing the inter-socket link.  Returns {"node", "cpus"} or None if nothing was done."""
    import os
    node = device_numa_node(device_index)
    if node is None or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            cpus = _parse_cpulist(fh.read())
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
    except (OSError, ValueError):
        return None
    return {"node": node, "cpus": len(cpus)}
