#!/usr/bin/env python
"""bench.py -- KeyPointDetection hot path (DoG pyramid + extrema/contrast/edge, Harris) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 1080p|4k|8k] [--frames F] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic frames: for every frame
vslam_dog_keypoints (frame -> ordered keypoint list) and vslam_harris in pipeline mode (frame -> response fp32 +
3x3-NMS mask + ordered NMS2 corner list).  value = input Mpix/s over all ranks with frames resident in HBM;
e2e = the same through the C ABI with HOST (pinned) frames and host result lists, copies inside the timed region.
Multi-GPU: frame-batch data parallelism, one process per GPU, no collective on the data path (weak scaling).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {"1080p": (1080, 1920, 256), "4k": (2160, 3840, 64), "8k": (4320, 7680, 16)}
MAC_PER_PX = 1738.0            # blur MACs per input pixel (both passes, 4 octaves x 6 levels)
IMMA_PEAK_TMAC = 568.9         # measured mma.sync.m16n8k32 u8 peak, TMAC/s (tools/imma_peak.cu, profiles/imma_peak_r1.json)
IDP_PEAK_TMAC = (74.31, 37.13)  # measured dp4a / dp2a peaks, TMAC/s (tools/alu_peak.cu, profiles/alu_peak_r1.json)
KP_CAP_PER_MPX = 40000         # output list capacity (entries per input Mpix), generous: ~1.2k/Mpx expected
METRIC = "Mpix/s (frames/s) DoG+Harris at 1/2/4/8 B200; % HBM roofline vs host CPU"


def bind_near_gpu(index):
    """One process per GPU: run (and first-touch the pinned host buffers) on the CPUs NVML reports as local to the GPU, so
    that eight ranks' uploads do not all cross the socket interconnect.  Returns the CPU count bound to, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([c.strip() for c in o.split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) >= 6 and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own CPU path = its OpenCV calls through cv2 (all host threads OpenCV
# uses) + its scalar C++ loops restated in plain C (oracle/vslam_oracle.c, single thread like the reference).
# ------------------------------------------------------------------------------------------------------------------
def cpu_reference_frames(frames):
    """Runs the CPU path on `frames`; returns (seconds, kind, cores, keypoints, corners)."""
    from oracle import c_oracle as O
    try:
        import cv2
        from oracle import cv2_reference as R
        have_cv2 = True
    except Exception:
        have_cv2 = False
    O.lib()
    t0 = time.perf_counter()
    nk = nc = 0
    for img in frames:
        if have_cv2:
            pyr = R.build_pyramid(img, 4, 1.6, gradients=True, mag_orient=True)      # GaussPyramid ctor, call for call
            for o in range(4):
                k, _ = O.detect_octave(pyr["dog"][o], pyr["gauss"][o], o, 3)       # scalar loops of Diff_of_Gauss.cpp
                nk += len(k)
            blurred, _, _ = R.harris_opencv_part(img)                               # Harris_corners.cpp:158-164
            L = O.lib()
            h, w = img.shape
            resp = np.empty((h, w), np.float32)
            L.vo_harris_response(O._u8(blurred), h, w, 0.04, O._f32(resp))          # scalar loops of Harris_corners.cpp
            r8 = np.empty((h, w), np.uint8); m3 = np.empty((h, w), np.uint8)
            L.vo_harris_nms3(O._f32(resp), h, w, O._u8(r8), O._u8(m3))
            cor = np.zeros((h * w // 4, 2), np.int32)
            nc += L.vo_harris_nms2(O._f32(resp), h, w, 5, cor.ctypes.data_as(O.i32p), None, len(cor), None)
        else:
            k, _, _ = O.dog_keypoints(img)
            nk += len(k)
            nc += len(O.harris(img)["corners"])
    dt = time.perf_counter() - t0
    cores = (cv2.getNumThreads() if have_cv2 else 1)
    kind = "port"   # the reference cannot be compiled here (needs OpenCV C++); cv2 runs its OpenCV calls, C restates its loops
    return dt, kind, cores, nk, nc


def run_reference(args, rank, world):
    if rank != 0:
        return
    from visualslam_b200 import synth
    h, w, _ = WORKLOADS[args.workload]
    n_sample = max(1, min(4, int(8e6 // (h * w)) or 1))       # bounded sample per step: ~8 Mpix
    frames = [synth.frame(i, h, w) for i in range(n_sample)]
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_reference_frames(frames[:1])
    times = []
    for _ in range(args.steps):
        dt, kind, cores, nk, nc = cpu_reference_frames(frames)
        times.append(dt)
    total = sum(times)
    mpix = n_sample * h * w * args.steps / 1e6 / total
    sample = f"{n_sample} synthetic {w}x{h} frames per step (ids 0..{n_sample - 1}) of the {args.workload} workload"
    line = {"impl": "reference", "metric": METRIC, "value": mpix, "unit": "Mpix/s", "frames_per_s": mpix * 1e6 / (h * w),
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"{args.workload}: DoG keypoints + Harris pipeline, bounded sample", "sample": sample},
            "cpu_baseline": {"value": mpix, "unit": "Mpix/s", "cores": cores, "kind": kind, "sample": sample,
                             "how": "cv2 (reference's OpenCV calls, OpenCV's own threads) + plain-C restatement of the reference's scalar loops (1 thread)"},
            "e2e": {"value": mpix, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="1080p", choices=list(WORKLOADS))
    ap.add_argument("--frames", type=int, default=0, help="frames per GPU per step (default: workload's)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-kernel-timing", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    import visualslam_b200 as vs
    from visualslam_b200 import synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    numa = bind_near_gpu(local) if world > 1 else None           # pinned buffers are first-touched on the GPU's own NUMA node
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    h, w, nf = WORKLOADS[args.workload]
    if args.frames:
        nf = args.frames
    px_step = nf * h * w
    stream = torch.cuda.Stream()
    ctx = vs.Context(local, stream=stream.cuda_stream)

    # ---- inputs: 8 unique frames cycled; device-resident copy + pinned host copy ----
    from visualslam_b200 import shard
    ids = shard.weak_batch(rank, world, nf)                      # this rank's frames of the global batch (contiguous block)
    uniq = [synth.frame(ids[0] + i, h, w) for i in range(min(8, nf))]
    host = torch.empty((nf, h, w), dtype=torch.uint8, pin_memory=True)
    hnp = host.numpy()
    for i in range(nf):
        hnp[i] = uniq[i % len(uniq)]
    dev = host.to(f"cuda:{local}")
    cap = int(KP_CAP_PER_MPX * px_step / 1e6) + 1024
    kp_dev = torch.zeros((cap, 6), dtype=torch.int32, device=dev.device)
    co_dev = torch.zeros((cap * 4, 3), dtype=torch.int32, device=dev.device)
    resp_dev = torch.empty((nf, h, w), dtype=torch.float32, device=dev.device)
    mask_dev = torch.empty((nf, h, w), dtype=torch.uint8, device=dev.device)

    def step_device():
        a = ctx.dog_keypoints(dev, out=kp_dev)
        b = ctx.harris(dev, device_out={"response": resp_dev, "mask3": mask_dev, "corners": co_dev})
        return sum(a["counts"]), sum(b["counts"]), a["overflow"] or b["overflow"]

    # e2e result lists: pinned host buffers owned by the caller (a pageable list would be staged by the driver)
    kp_host = torch.empty((cap, 6), dtype=torch.int32, pin_memory=True).numpy()
    co_host = torch.empty((cap * 4, 3), dtype=torch.int32, pin_memory=True).numpy()

    def step_e2e():
        up = ctx.upload(hnp)                                          # one chunked H2D of the batch, overlapped with the kernels
        a = ctx.dog_keypoints(up, out=kp_host)                        # keypoint list D2H inside
        b = ctx.harris(up, want_maps=False, corners_out=co_host)      # corner-only mode on the e2e path, corner list D2H inside
        return len(a["keypoints"]), len(b["corners"])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(steps):
                out = fn()
            e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    for _ in range(args.warmup):
        nk, nc, ovf = step_device()
    if ovf:
        raise SystemExit("output list overflow in warm-up")
    sampler = ClockSampler(local); sampler.start()
    l0 = ctx.launches
    if not args.no_kernel_timing:
        ctx.profile(True)
    ms, (nk, nc, ovf) = timed(step_device, args.steps)
    launches = ctx.launches - l0
    prof = ctx.profile_read() if not args.no_kernel_timing else {}
    ctx.profile(False)
    clocks = sampler.summary()
    ms_step = ms / args.steps
    value = world * px_step / 1e6 / (ms_step * 1e-3)

    e2e = None
    if not args.no_e2e:
        step_e2e()
        ems, (ek, ec) = timed(step_e2e, max(1, min(args.steps, 2)))
        ems /= max(1, min(args.steps, 2))
        e2e = {"value": world * px_step / 1e6 / (ems * 1e-3), "unit": "Mpix/s", "ms_per_step": ems,
               "h2d_bytes_per_step": px_step, "d2h_bytes_per_step": int(ek * 24 + ec * 12 + 8 * nf),
               "api": "vslam_upload_frames (chunked, overlapped) + vslam_dog_keypoints + vslam_harris(corner-only): host (pinned) frames in, host (pinned) result lists out"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    hbm_peak, peak_src = peaks()
    line = {"metric": METRIC, "value": value, "unit": "Mpix/s", "frames_per_s": value * 1e6 / (h * w), "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"synthetic {w}x{h} grayscale batch of {nf} frames per GPU, DoG (4 octaves x 6 levels, sigma 1.6) "
                                   f"+ Harris pipeline mode, {world} B200", "frames_per_gpu": nf, "rows": h, "cols": w,
                       "l2": f"inputs larger than L2 ({px_step / 1e6:.0f} MB batch per step, 8 unique frames cycled)",
                       "parallelism": f"frame-batch dp{world}, no data-path collective",
                       "host_cpus_bound": numa},
            "keypoints_per_step": int(nk), "corners_per_step": int(nc), "clocks": clocks, "gpu_launches": int(launches),
            "e2e": e2e}
    if prof:
        tot = sum(v["ms"] for v in prof.values()) or 1.0
        dom = max(prof, key=lambda k: prof[k]["ms"])
        dlaunch = prof[dom]["launches"] or 1
        # ALGORITHMIC HBM bytes per launch of each kernel, launches cover the whole batch:
        #   octave_fused_k (keypoints-only mode): read the octave base 1 B/px, write the next base 1/4 B/px (none after the
        #     last octave), 24 B per keypoint; 4 launches per step, one per octave -> average over the 4
        #   harris_fused_k (pipeline mode): read 1 B, write response 4 B + mask 1 B per pixel, 12 B per corner
        #   upsample2x_k: read 1 B, write 4 B per input pixel
        oct_px = [4.0 * h * w / 4 ** o for o in range(4)]
        fused_bytes = nf * (sum(oct_px) + sum(oct_px[1:])) / 4 + 24.0 * nk / 4
        per_launch = {"octave_fused_k": fused_bytes, "harris_fused_k": 6.0 * h * w * nf + 12.0 * nc,
                      "upsample2x_k": 5.0 * h * w * nf}
        alg = per_launch.get(dom, 0.0)
        ach = alg / 1e9 / (prof[dom]["ms"] / dlaunch * 1e-3) if alg else None
        fused_ms = prof["octave_fused_k"]["ms"]
        # The Gaussian stack (1738 MAC per input pixel) runs on the integer tensor pipe as banded-Toeplitz
        # products (IMMA.16832.U8.U8); useful MAC/s against the measured IMMA issue peak, and against what the packed
        # dot-product pipe could do at its measured peak (the bound of the dp4a/dp2a kernel this replaced)
        tmac = MAC_PER_PX * px_step * args.steps / (fused_ms * 1e-3) / 1e12 if fused_ms else None
        idp_bound = 2.0 / (1.0 / IDP_PEAK_TMAC[0] + 1.0 / IDP_PEAK_TMAC[1])   # half the MACs on dp4a, half on dp2a
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic_r1.json")
        if os.path.exists(tp):
            tj = json.load(open(tp))
            key = f"{args.workload}x{nf}"
            if key in tj and dom in tj[key]:
                traffic = tj[key][dom]
        line["roofline"] = {"bound": "hbm", "kernel": dom + (" (octave_mma_k<NCH>: tensor-pipe blur)" if dom == "octave_fused_k" else ""),
                            "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                            "frac": ach / hbm_peak if ach else None, "traffic": traffic, "peak_source": peak_src,
                            "mode": "keypoints-only (fused): octave bases are the only maps in HBM",
                            "note": "the fused octave kernel is not HBM bound (11.6 algorithmic bytes against 1738 MAC per input pixel, "
                                    "SURVEY.md 8(d)); the HBM fraction is reported because the contract asks for it, the tensor block says "
                                    "how far the blur is from the pipe it runs on; the kernel is issue/latency bound (profiles/)",
                            "tensor": {"pipe": "IMMA.16832.U8.U8 (mma.sync u8 x u8 -> s32), exact integer arithmetic",
                                       "useful_TMAC_s": tmac, "peak_TMAC_s": IMMA_PEAK_TMAC,
                                       "frac": tmac / IMMA_PEAK_TMAC if tmac else None,
                                       "vs_idp_pipe_bound": tmac / idp_bound if tmac else None,
                                       "idp_pipe_bound_TMAC_s": idp_bound,
                                       "peak_source": "measured on a B200 of this pool: tools/imma_peak.cu, tools/alu_peak.cu (profiles/*_peak_r1.json)",
                                       "mac_per_input_px": MAC_PER_PX},
                            "other_kernels_hbm": {k: {"achieved_GBps": per_launch[k] / 1e9 / (prof[k]["ms"] / max(prof[k]["launches"], 1) * 1e-3),
                                                      "frac": per_launch[k] / 1e9 / (prof[k]["ms"] / max(prof[k]["launches"], 1) * 1e-3) / hbm_peak}
                                                  for k in per_launch if k != dom and prof.get(k, {}).get("launches")},
                            "kernel_ms_per_step": {k: v["ms"] / args.steps for k, v in prof.items() if v["launches"]},
                            "kernel_launches_per_step": {k: v["launches"] / args.steps for k, v in prof.items() if v["launches"]},
                            "kernel_share": {k: v["ms"] / tot for k, v in prof.items() if v["launches"]}}
    if not args.no_cpu_baseline:
        n_sample = 8 if h * w <= 2.1e6 else (2 if h * w <= 8.3e6 else 1)     # 5-15 s of CPU work
        dt, kind, cores, _, _ = cpu_reference_frames([uniq[i] for i in range(min(n_sample, len(uniq)))])
        line["cpu_baseline"] = {"value": min(n_sample, len(uniq)) * h * w / 1e6 / dt, "unit": "Mpix/s", "cores": cores, "kind": kind,
                                "sample": f"{min(n_sample, len(uniq))} frames of the same workload ({dt:.1f} s of CPU work)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
