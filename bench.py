#!/usr/bin/env python
"""bench.py -- KeyPointDetection hot path (DoG pyramid + extrema/contrast/edge, Harris) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 1080p|4k|8k] [--frames F] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic frames: for every frame
vslam_dog_keypoints (frame -> ordered keypoint list) and vslam_harris in pipeline mode (frame -> response fp32 +
3x3-NMS mask + ordered NMS2 corner list).  value = input Mpix/s over all ranks with frames resident in HBM;
e2e = the same through the C ABI with HOST (pinned) frames and host result lists, copies inside the timed region.
Multi-GPU: frame-batch data parallelism, one process per GPU, no collective on the data path (weak scaling); before
the timed region every rank runs the SAME eight frames once and the ranks compare the md5 of their result lists.
The line also carries a `sweep` block (unless --no-sweep): the 4K and 8K batches of BASELINE.json's configs 4 and 5,
the single-frame 8K latency, and the materialise mode (everything the reference's GaussPyramid constructor produces).
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
MATERIALISE_ALL = 0x78          # VSLAM_MATERIALISE_ALL: gx, gy, magnitude, orientation of every level (include/vslam.h)
MATERIALISE_BYTES_PER_PX = 1 + 5.3125 * (11 + 96)   # 569.4 B per input pixel
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
    """(HBM GB/s, dense bf16 TFLOP/s sustained, source): the roofline denominators."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return j.get("hbm_gbs", 6650.0), j.get("bf16_tflops_sustained", j.get("bf16_tflops", 1590.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1400.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        # NVML in-process (a sample takes well under a millisecond, so even a 100 ms timed region is sampled many times: nvidia-smi
        # needs 0.1-0.3 s per call, longer with 8 ranks calling it at once); the nvidia-smi query of the recipe is the fallback
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            bits = [getattr(N, n, 0) for n in ("nvmlClocksEventReasonHwSlowdown", "nvmlClocksEventReasonHwThermalSlowdown",
                                                "nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksEventReasonSwPowerCap")]
            if not all(bits):
                bits = [getattr(N, n) for n in ("nvmlClocksThrottleReasonHwSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown",
                                                "nvmlClocksThrottleReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwPowerCap")]
            reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
            mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
            while not self.stop_flag:
                r = int(reasons(h))
                self.rows.append([str(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)), str(mx)] + ["Active" if r & b else "Not Active" for b in bits])
                time.sleep(0.005)
            return
        except Exception:
            pass
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
    # kind "port": cv2 (the real, threaded OpenCV) runs the reference's OpenCV calls and plain C restates its loops.  oracle/_ref -- the
    # reference's own translation units -- exists since round 2, but it is compiled against a stand-in for OpenCV whose imgproc is the
    # single-threaded C oracle: right for parity, far slower than OpenCV, so timing it would flatter the GPU.
    kind = "port"
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
            "config": {"workload": f"{args.workload}: DoG keypoints + Harris pipeline, bounded sample", "sample": sample,
                       "cpu_work": "the reference's stock path: the GaussPyramid constructor forms all 48 gradient / magnitude / phase maps per frame "
                                   "(GaussPyramid.cpp:65-104) that the GPU's keypoints-only path never forms (~2.7x of the CPU time, SURVEY.md 6)"},
            "cpu_baseline": {"value": mpix, "unit": "Mpix/s", "cores": cores, "kind": kind, "sample": sample,
                             "how": "cv2 (reference's OpenCV calls, OpenCV's own threads) + plain-C restatement of the reference's scalar loops (1 thread)"},
            "e2e": {"value": mpix, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------------------
class Runner:
    """One GPU's side of the bench: owns the context, runs a workload (h, w, frames) device-resident and end to end."""

    def __init__(self, args, rank, world, local):
        import torch
        import torch.distributed as dist

        import visualslam_b200 as vs
        self.torch, self.dist, self.vs = torch, dist, vs
        self.args, self.rank, self.world, self.local = args, rank, world, local
        self.device = torch.device("cuda", local)
        self.stream = torch.cuda.Stream()
        self.ctx = vs.Context(local, stream=self.stream.cuda_stream)
        m = getattr(args, "e2e_mode", "auto")
        self.e2e_mode = m if m != "auto" else ("stream" if world <= 4 else "sync")

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps):
        """CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        out = None
        with torch.cuda.stream(self.stream):
            e0.record(self.stream)
            for _ in range(steps):
                out = fn()
            e1.record(self.stream)
        self.barrier()
        ms = e0.elapsed_time(e1)
        if self.world > 1:
            t = torch.tensor([ms], device=self.device)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    def frames(self, h, w, nf, same_on_all_ranks=False):
        """8 unique synthetic frames cycled to fill the batch: pinned host copy + device copy."""
        from visualslam_b200 import shard, synth
        torch = self.torch
        first = 0 if same_on_all_ranks else shard.weak_batch(self.rank, self.world, nf)[0]
        uniq = [synth.frame(first + i, h, w) for i in range(min(8, nf))]
        host = torch.empty((nf, h, w), dtype=torch.uint8, pin_memory=True)
        hnp = host.numpy()
        for i in range(nf):
            hnp[i] = uniq[i % len(uniq)]
        return uniq, hnp, host.to(self.device)

    def workload(self, h, w, nf, steps, warmup, e2e=True, kernel_timing=True, clocks=False, e2e_sync=False):
        """Device-resident value + (optionally) the end-to-end figure of one (h, w, frames-per-GPU) workload."""
        torch, ctx = self.torch, self.ctx
        px_step = nf * h * w
        uniq, hnp, dev = self.frames(h, w, nf)
        cap = int(KP_CAP_PER_MPX * px_step / 1e6) + 1024
        kp_dev = torch.zeros((cap, 6), dtype=torch.int32, device=self.device)
        co_dev = torch.zeros((cap * 4, 3), dtype=torch.int32, device=self.device)
        resp_dev = torch.empty((nf, h, w), dtype=torch.float32, device=self.device)
        mask_dev = torch.empty((nf, h, w), dtype=torch.uint8, device=self.device)
        overflow = [False]

        def step_device():
            a = ctx.dog_keypoints(dev, out=kp_dev)
            b = ctx.harris(dev, device_out={"response": resp_dev, "mask3": mask_dev, "corners": co_dev})
            overflow[0] |= bool(a["overflow"] or b["overflow"])
            return sum(a["counts"]), sum(b["counts"])

        for _ in range(warmup):
            step_device()
        sampler = None
        if clocks:
            sampler = ClockSampler(self.local); sampler.start()
        l0 = ctx.launches
        if kernel_timing:
            ctx.profile(True)
        ms, (nk, nc) = self.timed(step_device, steps)
        launches = ctx.launches - l0
        prof = ctx.profile_read() if kernel_timing else {}
        ctx.profile(False)
        res = {"h": h, "w": w, "frames_per_gpu": nf, "px_step": px_step, "ms_per_step": ms / steps,
               "value": self.world * px_step / 1e6 / (ms / steps * 1e-3), "keypoints": nk, "corners": nc, "launches": launches,
               "prof": prof, "uniq": uniq, "clocks": sampler.summary() if sampler else None}
        if overflow[0]:
            raise SystemExit(f"output list overflow in the device-resident steps of {w}x{h} x {nf}")
        if e2e:
            # e2e result lists: pinned host buffers owned by the caller (a pageable list would be staged by the driver); two sets, because
            # a step's lists are completed (vslam_finish) after the next step has been enqueued
            kp_host = [torch.empty((cap, 6), dtype=torch.int32, pin_memory=True).numpy() for _ in range(2)]
            co_host = [torch.empty((cap * 4, 2), dtype=torch.int32, pin_memory=True).numpy() for _ in range(2)]   # vslam_corner8 entries (VSLAM_COMPACT)
            eo = [False]
            it = [0]
            pend = [ctx.upload(hnp)]                                          # batch 0 on its way; every step uploads the NEXT step's batch
            last = [(0, 0)]

            def step_e2e():
                # the stream-of-batches loop of INTEGRATION.md: enqueue both detectors on batch i (deferred), complete batch i-1 -- its
                # keypoint list and 8-byte corner list leave under batch i's kernels -- then start the upload of batch i+1.  Every step = one upload, one
                # detector call, one completed pair of host lists.
                i = it[0]; it[0] += 1
                ctx.dog_harris(pend[0], kp_out=kp_host[i & 1], co_out=co_host[i & 1], compact=True, defer=True)
                if i:
                    r = ctx.finish()                                          # batch i-1's lists: device-to-host first ...
                    eo[0] |= bool(r["overflow"])
                    last[0] = (len(r["keypoints"]), len(r["corners"]))
                pend[0] = ctx.upload(hnp)                                     # ... then batch i+1's upload, both under batch i's kernels (an
                return last[0]                                                # upload waits only for the readers of the buffer it overwrites)

            def step_e2e_sync():
                # latency regime: upload, both detectors, lists complete on return -- nothing overlaps across steps
                r = ctx.dog_harris(ctx.upload(hnp), kp_out=kp_host[0], co_out=co_host[0], compact=True)
                eo[0] |= bool(r["overflow"])
                return len(r["keypoints"]), len(r["corners"])

            def step_e2e_lookahead_sync():
                # many GPUs on one host: the synchronous call (lists drain chunk by chunk inside it) with the next batch's upload
                # started first spreads the copies over the step; measured better than the deferred loop at 8 GPUs
                up = pend[0]
                pend[0] = ctx.upload(hnp)
                r = ctx.dog_harris(up, kp_out=kp_host[0], co_out=co_host[0], compact=True)
                eo[0] |= bool(r["overflow"])
                return len(r["keypoints"]), len(r["corners"])

            mode = "latency" if e2e_sync else self.e2e_mode
            if e2e_sync:
                step_e2e = step_e2e_sync
            elif mode == "sync":
                step_e2e = step_e2e_lookahead_sync
            step_e2e()
            step_e2e()
            ems, (ek, ec) = self.timed(step_e2e, steps)
            r = ctx.finish()                                                  # the last step's lists (the step before the timed region was
            eo[0] |= bool(r and r["overflow"])                                # completed inside it: one completion per timed step)
            if eo[0]:
                raise SystemExit(f"output list overflow in the end-to-end steps of {w}x{h} x {nf}")
            ems /= steps
            res["e2e"] = {"value": self.world * px_step / 1e6 / (ems * 1e-3), "unit": "Mpix/s", "ms_per_step": ems, "steps": steps, "mode": mode,
                          "h2d_bytes_per_step": px_step, "d2h_bytes_per_step": int(ek * 24 + ec * 8 + 8 * nf),
                          "api": "vslam_upload_frames (chunked; the next step's batch is uploaded under this step's kernels: two ingest buffers) + vslam_dog_harris | VSLAM_DEFER + vslam_finish (= vslam_dog_keypoints + corner-only vslam_harris "
                                 "in one call; a step's lists are completed after the next step has been enqueued and leave under its kernels; VSLAM_COMPACT: 8-byte vslam_corner8 list, lossless): host (pinned) frames in, "
                                 "host (pinned) result lists out.  value's steps run Harris in pipeline mode (response + mask written to HBM, 6 B per pixel) and "
                                 "e2e's corner-only (nothing but the lists leaves the GPU), so e2e can exceed value by the 0.3 ms that costs"}
        del kp_dev, co_dev, resp_dev, mask_dev, dev
        torch.cuda.empty_cache()
        return res

    def materialise(self, h, w, nf, steps, warmup):
        """The reference constructor's real output (GaussPyramid.cpp:106-131: 6 Gaussians + 5 DoG u8, gx / gy / magnitude /
        orientation fp32 of every level, all octaves = 569.4 B per input pixel) + the detector on the materialised pyramid."""
        torch, ctx = self.torch, self.ctx
        _, _, dev = self.frames(h, w, nf)
        keep = {}

        pyr = ctx.build_pyramid(dev, 4, 1.6, materialise=MATERIALISE_ALL)      # storage once; every step rebuilds every map in place

        def step():
            ctx.rebuild_pyramid(pyr, dev)
            r = ctx.dog_detect(pyr, stage_counts=False)
            keep["n"] = sum(r["counts"]); keep["ovf"] = r["overflow"]
            return keep["n"]

        for _ in range(warmup):
            step()
        ctx.profile(True)
        ms, nk = self.timed(step, steps)
        prof = ctx.profile_read()
        ctx.profile(False)
        pyr.close()
        del dev
        torch.cuda.empty_cache()
        ms_step = ms / steps
        px = nf * h * w
        hbm, _, src = peaks()
        build_ms = sum(prof[k]["ms"] for k in ("upsample2x_k", "octave_mma_k", "octave_fused_k", "grad_k") if k in prof) / steps
        gbs = MATERIALISE_BYTES_PER_PX * px / 1e9 / (build_ms * 1e-3) if build_ms else None
        return {"workload": f"{w}x{h} x {nf} frames per GPU: vslam_pyramid_rebuild (VSLAM_MATERIALISE_ALL, in place) + vslam_dog_detect", "ms_per_step": ms_step,
                "value": self.world * px / 1e6 / (ms_step * 1e-3), "unit": "Mpix/s", "keypoints": int(nk), "overflow": bool(keep.get("ovf")),
                "bytes_per_input_px": MATERIALISE_BYTES_PER_PX,
                "build_kernels_ms_per_step": build_ms, "build_achieved_GBps": gbs, "build_hbm_frac": gbs / hbm if gbs else None, "peak_source": src,
                "kernel_ms_per_step": {k: v["ms"] / steps for k, v in prof.items() if v["launches"]},
                "note": "value is timed over the whole call incl. the host-side pyramid handle and the (pageable) keypoint list; the HBM fraction "
                        "uses the device time of the building kernels (upsample2x_k + octave_mma_k, materialising instantiation + grad_k) against 569.4 algorithmic bytes per input pixel"}

    def parity_across_ranks(self, h, w):
        """Un-timed: every rank runs the SAME 8 frames; the md5 of (keypoint list, corner list) must agree on all ranks."""
        import hashlib
        torch = self.torch
        _, hnp, dev = self.frames(h, w, 8, same_on_all_ranks=True)
        a = self.ctx.dog_keypoints(dev)
        b = self.ctx.harris(dev, want_maps=False, capacity=8 * h * w // 8)
        if a["overflow"] or b["overflow"]:
            raise SystemExit("output list overflow in the multi-GPU parity step")
        dig = hashlib.md5(np.ascontiguousarray(a["keypoints"]).tobytes() + np.ascontiguousarray(b["corners"]).tobytes()).digest()
        mine = torch.tensor(list(dig), dtype=torch.uint8, device=self.device)
        if self.world > 1:
            allv = [torch.empty_like(mine) for _ in range(self.world)]
            self.dist.all_gather(allv, mine)
            same = all(bool((v == allv[0]).all().item()) for v in allv)
        else:
            same = True
        del dev
        return same, dig.hex(), len(a["keypoints"]), len(b["corners"])


def measured_traffic(kernel, h, w, nf):
    """DRAM bytes per launch of `kernel` from the committed ncu capture of this workload (tools/prof_all.sh + tools/ncu_summary.py)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic_r2.json")))
        key = {(1080, 1920, 256): "1080px256"}.get((h, w, nf))
        return float(t[key][kernel]) if key and kernel in t.get(key, {}) else None
    except Exception:
        return None


def roofline_block(res, steps, world):
    """roofline of the dominant kernel of the headline workload + HBM fractions of the other kernels on the timed path."""
    prof, h, w, nf = res["prof"], res["h"], res["w"], res["frames_per_gpu"]
    px_step, nk, nc = res["px_step"], res["keypoints"], res["corners"]
    hbm_peak, bf16_peak, peak_src = peaks()
    live = {k: v for k, v in prof.items() if v["launches"]}
    tot = sum(v["ms"] for v in live.values()) or 1.0
    per_ms = {k: v["ms"] / max(v["launches"], 1) for k, v in live.items()}       # average launch duration
    # ALGORITHMIC HBM bytes per launch; the octave kernels: read the octave base 1 B/px, write the next base
    # 1/4 B/px (none after the last octave), 24 B per keypoint -- four launches per step, one per octave: the average
    oct_px = [4.0 * h * w / 4 ** o for o in range(4)]
    octave_bytes = nf * (sum(oct_px) + sum(oct_px[1:])) / 4 + 24.0 * nk / 4
    alg = {"octave_mma_k": octave_bytes, "octave_umma_k": octave_bytes, "octave_fused_k": octave_bytes,
           "harris_fast_k": 6.0 * h * w * nf + 12.0 * nc, "upsample2x_k": 5.0 * h * w * nf}
    octave_kernels = [k for k in ("octave_mma_k", "octave_umma_k", "octave_fused_k") if k in live]
    dom = max(live, key=lambda k: live[k]["ms"])
    blur_ms = sum(live[k]["ms"] for k in octave_kernels)
    # The Gaussian stack is 1738 u8 MAC per input pixel evaluated exactly on the integer tensor pipe; the kernel
    # is bound by the shared-memory traffic and the instructions around the MMAs, not by HBM: of the two rooflines the contract
    # offers, the tensor one is the binding one ("limiter" says what ncu shows).
    tmac = MAC_PER_PX * px_step * steps / (blur_ms * 1e-3) / 1e12 if blur_ms else None
    tops = 2.0 * tmac if tmac else None
    hbm_ach = alg.get(dom, 0.0) / 1e9 / (per_ms[dom] * 1e-3) if alg.get(dom) else None
    block = {"bound": "tensor" if dom in octave_kernels else "hbm", "kernel": dom,
             "achieved": tops if dom in octave_kernels else hbm_ach,
             "peak": bf16_peak if dom in octave_kernels else hbm_peak, "unit": "TFLOP/s" if dom in octave_kernels else "GB/s",
             "frac": (tops / bf16_peak if tops else None) if dom in octave_kernels else (hbm_ach / hbm_peak if hbm_ach else None),
             "traffic": measured_traffic(dom, h, w, nf), "peak_source": peak_src,
             "mode": "keypoints-only (fused): octave bases are the only maps in HBM",
             "limiter": "shared-memory pipe, not HBM and not the tensor pipe: l1tex__data_pipe_lsu_wavefronts is 65 % / 67 % of its peak at octaves "
                        "0 / 1 in the committed ncu capture of this command (profiles/ncu_full_r2_summary.csv; 68 % / 72 % before the V pass shared its "
                        "A fragments, DESIGN.md 4.1), issue slots 50 % busy, tensor pipe 25-30 %, DRAM traffic = algorithmic bytes",
             "note": "achieved = useful integer ops of the blur (2 x 1738 MAC per input pixel: the banded-Toeplitz zeros the MMAs also multiply are "
                     "not counted) over the device time of the octave kernels; peak = the measured dense bf16 rate (sustained) of MEASURED_PEAKS.json "
                     "-- the u8 rate of the tensor pipe is nominally twice that, so the fraction is an upper bound of the pipe's utilisation. "
                     "traffic: dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel (mean of the four octave launches of a "
                     "step) from the committed ncu --set full capture of the same command (profiles/traffic_r2.json, profiles/"
                     "ncu_full_r2_summary.csv); null for a workload that has no capture.  hbm.algorithmic_bytes_per_launch is the figure it is "
                     "compared with.",
             "tensor": {"useful_TMAC_s": tmac, "mma_sync_u8_peak_TMAC_s": IMMA_PEAK_TMAC, "frac_of_mma_sync_peak": tmac / IMMA_PEAK_TMAC if tmac else None,
                        "mac_per_input_px": MAC_PER_PX,
                        "peak_source": "mma.sync.m16n8k32 u8 issue peak measured on a B200 of this pool (tools/imma_peak.cu, profiles/imma_peak_r1.json)"},
             "hbm": {"achieved_GBps": hbm_ach, "frac": hbm_ach / hbm_peak if hbm_ach else None, "peak": hbm_peak,
                     "algorithmic_bytes_per_launch": alg.get(dom)},
             "other_kernels_hbm": {k: {"achieved_GBps": alg[k] / 1e9 / (per_ms[k] * 1e-3), "frac": alg[k] / 1e9 / (per_ms[k] * 1e-3) / hbm_peak}
                                   for k in alg if k in live and k != dom},
             "kernel_ms_per_step": {k: v["ms"] / steps for k, v in live.items()},
             "kernel_launches_per_step": {k: v["launches"] / steps for k, v in live.items()},
             "kernel_share": {k: v["ms"] / tot for k, v in live.items()}}
    return block


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
    ap.add_argument("--e2e-mode", default="auto", choices=["auto", "stream", "sync"],
                    help="end-to-end loop: stream = deferred calls (vslam_dog_harris | VSLAM_DEFER + vslam_finish, look-ahead uploads), "
                         "sync = one synchronous vslam_dog_harris per step with a look-ahead upload; auto = stream up to 4 GPUs per host, sync beyond")
    ap.add_argument("--no-sweep", action="store_true", help="skip the 4K / 8K / latency / materialise lines of the sweep block")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    from visualslam_b200 import shard
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    numa = bind_near_gpu(local) if world > 1 else None           # pinned buffers are first-touched on the GPU's own NUMA node
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    R = Runner(args, rank, world, local)
    h, w, nf = WORKLOADS[args.workload]
    if args.frames:
        nf = args.frames

    # ---- un-timed: the same 8 frames on every rank, result lists compared across ranks ----
    same, digest, pk, pc = R.parity_across_ranks(h, w)
    if not same:
        raise SystemExit("multi-GPU parity FAILED: ranks disagree on the result lists of the same frames")

    # ---- headline workload ----
    res = R.workload(h, w, nf, args.steps, max(args.warmup, 1), e2e=not args.no_e2e, kernel_timing=not args.no_kernel_timing, clocks=True)
    nk, nc, ms_max = shard.reduce_counts(res["keypoints"], res["corners"], res["ms_per_step"], device=R.device)

    # ---- sweep: BASELINE.json configs 4 and 5 per GPU, single-frame latency, materialise mode ----
    sweep = None
    if not args.no_sweep:
        sweep = {}
        ss, sw_ = max(2, min(args.steps, 3)), 1
        for name in ("4k", "8k"):
            if name == args.workload and not args.frames:
                continue
            sh, sw2, snf = WORKLOADS[name]
            r = R.workload(sh, sw2, snf, ss, sw_, e2e=not args.no_e2e, kernel_timing=False)
            sweep[f"{name}x{snf}"] = {"workload": f"synthetic {sw2}x{sh} x {snf} frames per GPU, DoG + Harris pipeline mode", "steps": ss,
                                      "value": r["value"], "unit": "Mpix/s", "frames_per_s": r["value"] * 1e6 / (sh * sw2), "ms_per_step": r["ms_per_step"],
                                      "e2e": r.get("e2e"), "keypoints_per_step": int(r["keypoints"]) * world, "corners_per_step": int(r["corners"]) * world}
        sh, sw2, _ = WORKLOADS["8k"]
        r = R.workload(sh, sw2, 1, max(args.steps, 5), 2, e2e=not args.no_e2e, kernel_timing=False, e2e_sync=True)
        sweep["8kx1_latency"] = {"workload": f"one synthetic {sw2}x{sh} frame per GPU per step (latency regime: 320 strips, one frame)",
                                 "ms_per_frame": r["ms_per_step"], "value": r["value"], "unit": "Mpix/s",
                                 "e2e_ms_per_frame": r["e2e"]["ms_per_step"] if r.get("e2e") else None,
                                 "e2e_mode": "synchronous call (upload + vslam_dog_harris, lists complete on return): a latency, not the stream-of-batches rate"}
        mh, mw, _ = WORKLOADS[args.workload]
        mnf = 8 if mh * mw <= 2.1e6 else (2 if mh * mw <= 8.3e6 else 1)           # 1.18 / 4.7 / 18.9 GB of maps per frame
        sweep["materialise"] = R.materialise(mh, mw, mnf, max(2, min(args.steps, 3)), 1)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    value = world * res["px_step"] / 1e6 / (ms_max * 1e-3)
    line = {"metric": METRIC, "value": value, "unit": "Mpix/s", "frames_per_s": value * 1e6 / (h * w), "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"synthetic {w}x{h} grayscale batch of {nf} frames per GPU, DoG (4 octaves x 6 levels, sigma 1.6) "
                                   f"+ Harris pipeline mode, {world} B200", "frames_per_gpu": nf, "rows": h, "cols": w,
                       "l2": f"inputs larger than L2 ({res['px_step'] / 1e6:.0f} MB batch per step, 8 unique frames cycled)",
                       "parallelism": f"frame-batch dp{world}, no data-path collective",
                       "host_cpus_bound": numa,
                       "cpu_arm": "the CPU arm (cpu_baseline / --impl reference) runs the reference's stock path, whose GaussPyramid constructor also forms "
                                  "the 48 gradient / magnitude / phase maps per frame (GaussPyramid.cpp:65-104); the GPU's keypoints-only path does not "
                                  "form them (~2.7x of the CPU time, SURVEY.md 6) -- sweep.materialise is the GPU mode that does"},
            "keypoints_per_step": int(nk), "corners_per_step": int(nc), "clocks": res["clocks"], "gpu_launches": int(res["launches"]),
            "multi_gpu_parity": bool(same), "parity_step": {"frames": 8, "md5": digest, "keypoints": pk, "corners": pc, "ranks_compared": world},
            "e2e": res.get("e2e")}
    if res["prof"]:
        line["roofline"] = roofline_block(res, args.steps, world)
    if sweep is not None:
        line["sweep"] = sweep
    if not args.no_cpu_baseline:
        uniq = res["uniq"]
        n_sample = 8 if h * w <= 2.1e6 else (2 if h * w <= 8.3e6 else 1)     # 5-15 s of CPU work
        dt, kind, cores, _, _ = cpu_reference_frames([uniq[i] for i in range(min(n_sample, len(uniq)))])
        line["cpu_baseline"] = {"value": min(n_sample, len(uniq)) * h * w / 1e6 / dt, "unit": "Mpix/s", "cores": cores, "kind": kind,
                                "sample": f"{min(n_sample, len(uniq))} frames of the same workload ({dt:.1f} s of CPU work)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
