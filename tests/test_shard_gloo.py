"""N > 1 host logic on CPU: two gloo ranks partition a batch by frames, each runs the (oracle) detector on its frames,
and the epilogue reduction gives the single-process totals and the slowest rank's time.  The GPU path uses exactly this
partition and reduction (bench.py), with the CUDA library in place of the oracle and NCCL in place of gloo."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from visualslam_b200 import shard, synth


def test_frame_range_is_a_partition():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 7, 8, 64, 1024, 1025):
            parts = [shard.frame_range(r, world, n) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    assert list(shard.weak_batch(1, 2, 4)) == [4, 5, 6, 7]
    with pytest.raises(ValueError):
        shard.frame_range(2, 2, 8)


def _worker(rank, world, port, n_frames, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import c_oracle as O
    lo, hi = shard.frame_range(rank, world, n_frames)
    nk = nc = 0
    for f in range(lo, hi):
        img = synth.frame(f, 96, 128)
        nk += len(O.dog_keypoints(img)[0])
        nc += len(O.harris(img)["corners"])
    tk, tc, tm = shard.reduce_counts(nk, nc, 10.0 + rank)
    if rank == 0:
        out.put((tk, tc, tm))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_partition_matches_single_process():
    from oracle import c_oracle as O
    O.lib()
    n_frames = 5
    want_k = sum(len(O.dog_keypoints(synth.frame(f, 96, 128))[0]) for f in range(n_frames))
    want_c = sum(len(O.harris(synth.frame(f, 96, 128))["corners"]) for f in range(n_frames))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got == (want_k, want_c, 11.0)
