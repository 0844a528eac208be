"""Frame-batch data parallelism across the GPUs of one box.

Every frame is processed independently (the reference handles exactly one image per run, Diff_of_Gauss.cpp:730,
Harris_corners.cpp:148), so the path shards by frames: one process per GPU, contiguous block partition of the batch,
NO collective on the data path.  The only cross-rank traffic is the optional epilogue below: a sum of two int64
(keypoints, corners) and the max of the step time -- 16 bytes over NCCL (or gloo in the CPU tests).
"""
import os


def env_rank_world():
    """(rank, local_rank, world) from the torchrun environment; (0, 0, 1) when launched plainly."""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def frame_range(rank, world, n_frames):
    """Contiguous block partition: rank g owns frames [g*B/G, (g+1)*B/G) of a batch of B = n_frames."""
    if world <= 0 or not 0 <= rank < world or n_frames < 0:
        raise ValueError(f"rank {rank} of {world}, {n_frames} frames")
    return (rank * n_frames) // world, ((rank + 1) * n_frames) // world


def weak_batch(rank, world, frames_per_gpu):
    """Weak scaling (bench.py): every rank processes frames_per_gpu frames; returns its global frame ids."""
    lo, hi = frame_range(rank, world, world * frames_per_gpu)
    return range(lo, hi)


def reduce_counts(keypoints, corners, ms, device=None):
    """Optional epilogue: whole-job totals and the slowest rank's time.  No-op without an initialised process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return int(keypoints), int(corners), float(ms)
    t = torch.tensor([int(keypoints), int(corners)], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    m = torch.tensor([float(ms)], dtype=torch.float64, device=device)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return int(t[0].item()), int(t[1].item()), float(m.item())
