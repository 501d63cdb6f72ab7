#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native LibGaussianBlur hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Metric (BASELINE.json): Mpixel/s on a 7680x4320 RGBA uint8 image, sigma=20, all four channels
blurred. A step = one full blur (row pass + column pass) of one synthetic image per GPU. With
N > 1 (torchrun, one rank per GPU) every rank blurs its own frame — frames shard with no
collective — so scaling is weak and `value` is the whole-job pixels / max-over-ranks time.

Prints ONE JSON line. `value` is device-resident throughput (CUDA events on the launch stream);
`e2e` goes through the host-buffer C-ABI call (pinned host memory, H2D + blur + D2H inside the
timed region); `roofline` is the dominant kernel against the measured HBM copy peak;
`cpu_baseline` is the compiled, unmodified reference (oracle/_ref) on the box's host cores.
`--impl reference` times only that CPU reference and prints the same shape of line.
"""
from __future__ import annotations

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

WORKLOADS = {
    # name: (rows, cols, channels, sigma, apply_to_alpha)   one image per GPU per step (weak scaling)
    "8k_rgba_sigma20": (4320, 7680, 4, 20.0, True),
    "1080p_rgb_sigma5": (1080, 1920, 3, 5.0, False),
    "1080p_rgb_sigma10": (1080, 1920, 3, 10.0, False),
    # BASELINE.json config 3: 512 frames sharded by frame over the ranks, no collective (strong scaling)
    "video_1080p_x512_sigma10": (1080, 1920, 3, 10.0, False),
    # BASELINE.json config 4: one 32768^2 single-channel image, row bands + all-to-all (strong scaling)
    "band_32k_gray_sigma50": (32768, 32768, 1, 50.0, False),
    # BASELINE.json config 5: 65536 x 32768 RGB, sigma=150 (FFT lengths 69120 / 34560), row bands + all-to-all
    "band_64kx32k_rgb_sigma150": (32768, 65536, 3, 150.0, False),
    # a 4x smaller band-sharded case for quick checks
    "band_16k_gray_sigma50": (16384, 16384, 1, 50.0, False),
}
VIDEO_FRAMES = 512
METRIC = "Mpixel/s 8K RGBA sigma=20"
UNIT = "Mpixel/s"


def load_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed ncu --set full capture of
    this workload (profiles/r01/traffic_8k.json; tools/gpu_final.sh regenerates it). None when there is no capture."""
    if workload != "8k_rgba_sigma20":
        return None, None
    p = os.path.join(ROOT, "profiles", "r01", "traffic_8k.json")
    try:
        t = json.load(open(p))[kernel]
        return int(t["traffic"]), t["source"]
    except Exception:
        return None, None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clocks and throttle reasons sampled DURING the timed region (NVML in a thread, 5 ms period;
    falls back to one `nvidia-smi` query)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.sm, self.reasons, self.power = [], set(), []
        self.smax = None
        self._stop = threading.Event()
        self.thread = None
        self.nvml = None

    def start(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nvml = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = self.idx
            if vis:
                try:
                    phys = int(vis.split(",")[self.idx])
                except ValueError:
                    phys = self.idx
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None
            return
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()

    def _loop(self):
        n = self.nvml
        while not self._stop.is_set():
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
                mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
                self.power.append(n.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        if self.nvml is None:
            try:
                out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits",
                                      "-i", str(self.idx)], capture_output=True, text=True, timeout=20).stdout
                a, b = [float(x) for x in out.strip().split(",")]
                return {"sm_mhz": a, "sm_max_mhz": b, "samples": 1, "reasons": [], "note": "single idle sample"}
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["clock query unavailable"]}
        self._stop.set()
        self.thread.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.smax,
                "samples": len(self.sm), "power_w_max": max(self.power) if self.power else None,
                "reasons": sorted(self.reasons)}


def cpu_reference_run(rows, cols, ch, sigma, alpha, runs, warm=1):
    """Times gaussianblur::gaussianblur of the compiled reference on all host cores."""
    from libgaussianblur_b200.synth import synth_u8
    from oracle import ref_lib  # the compiled, unmodified reference: the thing this leg measures

    if not ref_lib.available():
        return None
    img = synth_u8((rows, cols, ch), 0x1234 + 1)
    times = []
    for i in range(warm + runs):
        _, secs = ref_lib.gaussianblur(img, sigma, alpha, return_seconds=True)
        if i >= warm:
            times.append(secs)
    return times, ref_lib.hw_threads()


def run_reference_arm(args, wl):
    rows, cols, ch, sigma, alpha = WORKLOADS[wl]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    res = cpu_reference_run(rows, cols, ch, sigma, alpha, runs=args.steps, warm=args.warmup)
    if res is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libgblur_ref.so was not built"}))
        return 0
    times, cores = res
    total = float(sum(times))
    mpx = rows * cols * len(times) / total / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": mpx, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl, "rows": rows, "cols": cols, "channels": ch, "sigma": sigma,
                   "apply_to_alpha": alpha, "timing": "steady_clock around gaussianblur::gaussianblur only"},
        "cpu_baseline": {"value": mpx, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": f"{len(times)} full {cols}x{rows}x{ch} blurs, all host threads"},
        "e2e": {"value": mpx, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def _events(torch):
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def bench_video(args, wl, torch, dist, gb, world, rank, local, barrier):
    """Config 3: VIDEO_FRAMES frames of 1080p RGB sharded by frame (contiguous blocks), one plan per rank,
    no collective. value = all frames' pixels / max-over-ranks time."""
    from libgaussianblur_b200.multi import FrameShardedBlur
    from libgaussianblur_b200.synth import synth_u8

    rows, cols, ch, sigma, alpha = WORKLOADS[wl]
    fs = FrameShardedBlur(VIDEO_FRAMES, rows, cols, ch, sigma, device=local, chunk_frames=args.chunk_frames)
    f0, f1 = fs.frames
    n_local = f1 - f0
    one = synth_u8((rows, cols, ch), 0x1234 + 2)
    d_in = torch.from_numpy(one).cuda().unsqueeze(0).repeat(n_local, 1, 1, 1).contiguous()
    d_in[:, 0, 0, 0] = torch.arange(f0, f1, device="cuda").to(torch.uint8)  # frames differ
    d_out = torch.empty_like(d_in)
    h_pin = torch.empty((n_local, rows, cols, ch), dtype=torch.uint8).pin_memory()
    h_pin.copy_(d_in.cpu())
    stream = torch.cuda.current_stream()
    for _ in range(max(args.warmup, 3)):
        fs.run(d_in, d_out, alpha)
    barrier()
    l0 = gb.launch_count()
    e0, e1 = _events(torch)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        fs.run(d_in, d_out, alpha)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = gb.launch_count() - l0
    # e2e: the rank's whole frame block goes host -> device -> host through ONE host-buffer C-ABI call, which
    # streams it in chunks of chunk_frames (H2D || blur || D2H)
    fs.plan.run_host(h_pin, alpha, n_frames=n_local)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fs.plan.run_host(h_pin, alpha, n_frames=n_local)
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    fs.close()
    npx_step = VIDEO_FRAMES * rows * cols
    cp = ch
    return dict(ms_total=ms, e2e_ms=e2e_ms, npx_step=npx_step, launches=launches, scaling="strong",
                alg_bytes_step=npx_step * 10 * cp, h2d=n_local * rows * cols * ch,
                d2h=n_local * rows * cols * ch,
                config={"workload": wl, "frames": VIDEO_FRAMES, "rows": rows, "cols": cols, "channels": ch, "sigma": sigma,
                        "apply_to_alpha": alpha, "parallelism": f"frames sharded x{world}, no collective",
                        "frames_per_rank": n_local, "chunk_frames": fs.chunk,
                        "l2": f"{n_local * rows * cols * ch / 1e6:.0f} MB u8 per rank per step, larger than L2"},
                phases_ms=None)


def bench_band(args, wl, torch, dist, gb, world, rank, local, barrier):
    """Config 4: one big single-channel image. 1 GPU: the fused run. N GPUs: row bands, row pass local,
    NCCL all-to-all of the fp32 intermediate, column pass local (output stays column-band-owned)."""
    from libgaussianblur_b200.multi import BandShardedBlur, shard_range
    from libgaussianblur_b200.synth import synth_u8

    rows, cols, ch, sigma, alpha = WORKLOADS[wl]
    r0, r1 = shard_range(rows, world, rank)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(0x1234 + 3 + rank)  # uniform u8 generated on the device (the bands are GBs)
    band = torch.randint(0, 256, (r1 - r0, cols, ch), dtype=torch.uint8, device="cuda", generator=gen)
    h_pin = torch.empty_like(band, device="cpu").pin_memory()
    h_pin.copy_(band.cpu())
    stream = torch.cuda.current_stream()
    phases = {}
    single = world == 1 and not args.force_band
    if single:
        plan = gb.Plan(rows, cols, ch, sigma, device=local)
        d_out = torch.empty_like(band)
        step = lambda: plan.run_device(band, d_out, alpha)
        e2e_step = lambda: plan.run_host(h_pin, alpha)
        h2d = d2h = band.numel()
        closer = plan.close
    else:
        bs = BandShardedBlur.cuda(rows, cols, ch, sigma, alpha, device=local, fused=(args.exchange if args.exchange in ("staged", "hybrid") else args.exchange == "fused"),
                                  pipeline_chunks=(args.pipeline_chunks if args.exchange in ("nccl_pipelined", "staged", "hybrid") else 0))
        step = lambda: bs.run(band)
        h_out = torch.empty((rows, bs.my_cols[1] - bs.my_cols[0], ch), dtype=torch.uint8).pin_memory()

        def e2e_step():
            d = h_pin.cuda(non_blocking=True)
            h_out.copy_(bs.run(d), non_blocking=True)
            torch.cuda.synchronize()

        h2d, d2h = band.numel(), h_out.numel()
        closer = bs.close
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    l0 = gb.launch_count()
    e0, e1 = _events(torch)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = gb.launch_count() - l0
    if single:
        for name, st in (("row_pass", 1), ("col_pass", 2)):
            k0, k1 = _events(torch)
            torch.cuda.synchronize()
            k0.record(stream)
            for _ in range(args.steps):
                plan.run_stage(st, band, d_out, alpha)
            k1.record(stream)
            torch.cuda.synchronize()
            phases[name] = k0.elapsed_time(k1) / args.steps
    else:
        if bs.fused_run is not None and args.exchange in ("fused", "staged", "hybrid"):
            bs.profile = True
            step()
            torch.cuda.synchronize()
            phases.update(bs.phase_ms())
            bs.profile = False
        phases["exchange"] = bs.exchange_kind
        phases["exchanged_bytes_per_rank"] = (r1 - r0) * (cols - cols // world) * ch * 4
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    closer()
    npx_step = rows * cols
    return dict(ms_total=ms, e2e_ms=e2e_ms, npx_step=npx_step, launches=launches, scaling="strong",
                alg_bytes_step=npx_step * 10 * ch, h2d=h2d, d2h=d2h,
                config={"workload": wl, "rows": rows, "cols": cols, "channels": ch, "sigma": sigma,
                        "apply_to_alpha": alpha,
                        "parallelism": "single GPU" if single else f"row bands x{world} + {args.exchange} exchange + column bands",
                        "l2": "inputs larger than L2"},
                phases_ms=phases)


def finish_multi(args, res, torch, dist, world, rank, sampler):
    t = torch.tensor([res["ms_total"], res["e2e_ms"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = float(t[0]), float(t[1])
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    peak, peak_src = load_peaks()
    per_step_s = ms_total / args.steps * 1e-3
    value = res["npx_step"] / per_step_s / 1e6
    achieved = res["alg_bytes_step"] / world / per_step_s / 1e9  # per GPU, whole step
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": res["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": res["config"],
        "e2e": {"value": res["npx_step"] / (e2e_ms / args.steps * 1e-3) / 1e6, "unit": UNIT,
                "h2d_bytes_per_step": int(res["h2d"]), "d2h_bytes_per_step": int(res["d2h"]),
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(res["launches"]),
        "roofline": {"bound": "hbm", "kernel": "whole step (row pass + column pass" + (" + exchange)" if world > 1 and "band" in res["config"]["workload"] else ")"),
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": int(res["alg_bytes_step"] // world)},
        "phases_ms": res["phases_ms"], "clocks": clocks,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="8k_rgba_sigma20", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--chunk-frames", type=int, default=16, help="frames per launch in the video workload")
    ap.add_argument("--force-band", action="store_true", help="band workloads at 1 GPU: run the band-sharded path anyway")
    ap.add_argument("--pipeline-chunks", type=int, default=8)
    ap.add_argument("--exchange", default="staged", choices=["fused", "staged", "hybrid", "nccl", "nccl_pipelined"],
                    help="band workloads: row pass stores into peer memory (fused), local row pass in chunks with "
                         "copy-engine pushes into the owners' panel buffers (staged), or NCCL all-to-all")
    args = ap.parse_args()
    wl = args.workload
    if args.impl == "reference":
        return run_reference_arm(args, wl)

    # rank 0 prints ONE JSON line: libraries (NCCL's version banner) write to fd 1, so park the real
    # stdout and point fd 1 at stderr until the line is ready
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout

    import torch
    import torch.distributed as dist

    import libgaussianblur_b200 as gb
    from libgaussianblur_b200.synth import synth_u8

    rows, cols, ch, sigma, alpha = WORKLOADS[wl]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    gb.lib()
    if wl.startswith("video_") or wl.startswith("band_"):
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        fn = bench_video if wl.startswith("video_") else bench_band
        res = fn(args, wl, torch, dist, gb, world, rank, local, barrier)
        finish_multi(args, res, torch, dist, world, rank, sampler)
        if world > 1:
            dist.destroy_process_group()
        return 0
    plan = gb.Plan(rows, cols, ch, sigma, device=local)
    npx = rows * cols
    img = synth_u8((rows, cols, ch), 0x1234 + 1 + rank)
    h_pinned = torch.from_numpy(img.copy()).pin_memory()
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    stream = torch.cuda.current_stream()

    # ---- device-resident throughput ----
    for _ in range(max(args.warmup, 3)):
        plan.run_device(d_in, d_out, alpha)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = gb.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        plan.run_device(d_in, d_out, alpha)
    e1.record(stream)
    barrier()
    launches = gb.launch_count() - launches0
    ms_total = e0.elapsed_time(e1)
    pass_kernels = plan.last_kernels()

    # ---- per-kernel durations (same stream, CUDA events) for the roofline ----
    cp = plan.processed_channels(alpha)
    kern_ms = {}
    for name, fn in (("row_pass", lambda: plan.run_stage(1, d_in, d_out, alpha)),
                     ("col_pass", lambda: plan.run_stage(2, d_in, d_out, alpha))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record(stream)
        for _ in range(args.steps):
            fn()
        k1.record(stream)
        torch.cuda.synchronize()
        kern_ms[name] = k0.elapsed_time(k1) / args.steps
    clocks = sampler.stop() if rank == 0 else None

    # ---- end to end through the host-buffer C-ABI call (pinned memory) ----
    for _ in range(2):
        plan.run_host(h_pinned, alpha)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        plan.run_host(h_pinned, alpha)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0

    t = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = float(t[0]), float(t[1])

    if rank == 0:
        peak, peak_src = load_peaks()
        value = world * npx * args.steps / (ms_total * 1e-3) / 1e6
        e2e_value = world * npx * args.steps / (e2e_ms * 1e-3) / 1e6
        dom = max(kern_ms, key=kern_ms.get)
        alg_bytes = npx * cp * 5 + (npx * (ch - cp) if dom == "col_pass" else 0)  # u8 + fp32 per pixel-channel
        achieved = alg_bytes / (kern_ms[dom] * 1e-3) / 1e9
        traffic, traffic_src = load_traffic(wl, dom)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl, "rows": rows, "cols": cols, "channels": ch, "sigma": sigma,
                       "apply_to_alpha": alpha, "frames_per_gpu_per_step": 1, "parallelism": f"frames x{world}",
                       "l2": "inputs larger than L2 (133 MB u8 + 531 MB fp32 intermediate per step)",
                       "fft_lengths": [plan.n_rowpass, plan.n_colpass],
                       "radices": [plan.radices(0)[0], plan.radices(1)[0]],
                       "row_pass_kernel": pass_kernels[0], "col_pass_kernel": pass_kernels[1]},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(img.nbytes),
                    "d2h_bytes_per_step": int(img.nbytes), "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": int(launches),
            "kernels_ms": kern_ms,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(alg_bytes),
                         "whole_step_frac": (npx * (10 * cp + 2 * (ch - cp)) / (ms_total / args.steps * 1e-3) / 1e9) / peak},
            "clocks": clocks,
        }
        if not args.no_cpu_baseline:
            res = cpu_reference_run(rows, cols, ch, sigma, alpha, runs=3, warm=1)
            if res is not None:
                times, cores = res
                line["cpu_baseline"] = {"value": npx * len(times) / sum(times) / 1e6, "unit": UNIT, "cores": cores,
                                        "kind": "reference",
                                        "sample": f"{len(times)} full {cols}x{rows}x{ch} blurs after 1 warm-up, all host threads"}
            else:
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                        "sample": "oracle/_ref not built"}
        print(json.dumps(line), flush=True)
    plan.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
