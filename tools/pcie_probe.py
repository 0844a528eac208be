#!/usr/bin/env python
"""Host<->device copy bandwidth per rank, alone and with every rank copying at once (name the limiter of
the end-to-end line at 8 GPUs).  Launch like bench.py:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/pcie_probe.py
Each rank binds to the CPUs NVML reports as local to its GPU (as bench.py does), pins 512 MiB and times, with CUDA events:
H2D alone (ranks take turns), D2H alone, H2D all ranks together, D2H together, both directions together.
Rank 0 prints one JSON line."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from bench import bind_near_gpu

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
bound = bind_near_gpu(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N = 512 << 20
h_in = torch.empty(N, dtype=torch.uint8, pin_memory=True); h_in.fill_(rank + 1)
h_out = torch.empty(N, dtype=torch.uint8, pin_memory=True); h_out.fill_(0)
d_a = torch.empty(N, dtype=torch.uint8, device="cuda"); d_b = torch.ones(N, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(h2d, d2h, reps=4):
    """GB/s of this rank: (h2d rate, d2h rate); the two directions run on two streams"""
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    with torch.cuda.stream(s1):
        e[0].record()
        for _ in range(reps if h2d else 0):
            d_a.copy_(h_in, non_blocking=True)
        e[1].record()
    with torch.cuda.stream(s2):
        e[2].record()
        for _ in range(reps if d2h else 0):
            h_out.copy_(d_b, non_blocking=True)
        e[3].record()
    torch.cuda.synchronize()
    up = reps * N / 1e9 / (e[0].elapsed_time(e[1]) * 1e-3) if h2d else 0.0
    dn = reps * N / 1e9 / (e[2].elapsed_time(e[3]) * 1e-3) if d2h else 0.0
    return up, dn


timed(True, True, 1)
res = {}
# alone: ranks take turns
for name, (u, d) in (("h2d_alone", (True, False)), ("d2h_alone", (False, True)), ("both_alone", (True, True))):
    mine = None
    for r in range(world):
        barrier()
        if r == rank:
            mine = timed(u, d)
    res[name] = mine
for name, (u, d) in (("h2d_together", (True, False)), ("d2h_together", (False, True)), ("both_together", (True, True))):
    barrier()
    res[name] = timed(u, d)
barrier()
t = torch.tensor([v for k in sorted(res) for v in res[k]], dtype=torch.float64, device="cuda")
if world > 1:
    allr = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allr, t)
else:
    allr = [t]
if rank == 0:
    keys = sorted(res)
    out = {"n_gpus": world, "bytes_per_copy": N, "cpus_bound": bound, "unit": "GB/s per rank: [h2d, d2h]"}
    for i, k in enumerate(keys):
        out[k] = [[round(float(a[2 * i]), 1), round(float(a[2 * i + 1]), 1)] for a in allr]
        out[k + "_sum"] = [round(sum(float(a[2 * i]) for a in allr), 1), round(sum(float(a[2 * i + 1]) for a in allr), 1)]
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
