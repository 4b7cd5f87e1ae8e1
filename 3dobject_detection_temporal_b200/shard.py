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
