#!/usr/bin/env python
"""bench.py — spectrum frames/s on BASELINE.json's headline configuration (config 3):
1024 stereo streams, 8192-point Blackman-Harris, Separate, Lanczos bin interpolation, 1024
log-spaced display bins, dB scaling + two peak-hold graphs, 64 frames per stream at 75 % overlap.

    python bench.py --gpus N --steps K --warmup W           # this repo's CUDA path
    python bench.py --impl reference --gpus N ...           # the reference's CPU algorithm (oracle)

A step = one pass of the hot path over one batch: every stream's 64 frames
(prepareTransform -> doTransform -> mapToLinearSpace -> postProcessStdTransform each).  Streams
are independent, so with N GPUs every rank runs its own 1024 streams (weak scaling, no data-path
collective); `value` is the whole job's frames divided by the slowest rank's device time."""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

ALGO_BYTES_PER_FRAME = 2 * 8192 * 4 + 2 * 2 * 1024 * 8 + 2 * 1024 * 8  # 114 688 B, SURVEY.md §8(d)
UNIQUE_BYTES_PER_FRAME = 2 * 2048 * 4 + 2 * 1024 * 8                    # new hop + results (state amortised)
METRIC = "spectrum frames/s (8192-pt, 1024 streams)"
UNIT = "frames/s"


def workload_config(args):
    return {
        "workload": "C3: batched Spectrum FFT, 1024 stereo streams x 8192-pt Blackman-Harris, Separate, Lanczos, "
                    "1024 log bins, dB + 2 peak-hold graphs",
        "streams_per_gpu": args.streams, "frames_per_stream": args.frames, "hop": args.hop,
        "transform_size": 8192, "axis_points": 1024,
        "frames_per_step_per_gpu": args.streams * args.frames,
        "l2_policy": "inputs larger than L2 (1.12 GB audio + 1.07 GB results per step per GPU)",
        "sharding": "independent streams per rank, no collective",
    }


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(power), "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes per launch of the fused kernel from the committed ncu summary, if any."""
    path = os.path.join(ROOT, "profiles", "fused_kernel_summary.json")
    try:
        with open(path) as f:
            return json.load(f).get("dram_bytes_per_launch")
    except Exception:
        return None


def bench_cfg():
    from signalizer_b200 import _abi
    return _abi.default_config()  # C3


def make_audio(args, rank):
    from signalizer_b200 import signals
    n = signals.frames_length(8192, args.hop, args.frames)
    # synthetic sines / chirps / noise; 96 distinct streams tiled to the batch keeps
    # host generation time bounded without changing the arithmetic
    base = signals.batch(min(args.streams, 96), n, first_stream=rank * 96)
    reps = -(-args.streams // base.shape[0])
    return np.ascontiguousarray(np.tile(base, (reps, 1, 1))[:args.streams]), n


def cpu_sample(args, threads, streams, frames, fast=True):
    """Time the oracle (the reference's CPU algorithm restated) on a bounded sample."""
    from oracle.binding import OracleSpectrum
    from signalizer_b200 import signals
    cfg = bench_cfg()
    n = signals.frames_length(8192, args.hop, frames)
    audio = signals.batch(streams, n)
    o = OracleSpectrum(cfg, streams, fast=fast)
    o.process_frames(audio[:, :8192], args.hop, 1, threads=threads, want_results=False)  # touch
    t0 = time.perf_counter()
    o.process_frames(audio, args.hop, frames, threads=threads, want_results=False)
    dt = time.perf_counter() - t0
    return streams * frames / dt, dt


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path cannot be built
    (External/cpl is absent), so this times the oracle port with every host thread."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import binding
    binding.build()
    cores = binding.hardware_concurrency() or os.cpu_count() or 1
    # a step large enough that thread start-up does not count against the reference (about 0.1 s of work on 16 cores),
    # small enough that the default --steps 50 ends within seconds
    streams = max(cores * 16, 64)
    frames = 16
    vals = []
    for i in range(args.warmup + args.steps):
        v, dt = cpu_sample(args, cores, streams, frames)
        if i >= args.warmup:
            vals.append((v, dt))
    value = statistics.median(v for v, _ in vals)
    ms = statistics.median(dt for _, dt in vals) * 1e3
    sample = f"{streams} streams x {frames} frames per step (same C3 config and signals), {cores} threads over streams"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "note": "restated reference algorithm (scalar Stockham FFT), not the shipped binary (pffft); parity unpinned"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from signalizer_b200 import _abi
    from signalizer_b200.spectrum import TransformPairBatch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bind_to_gpu_numa_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cfg = bench_cfg()
    audio_np, n = make_audio(args, rank)
    S, F, hop, P = args.streams, args.frames, args.hop, cfg.axis_points
    pairs = TransformPairBatch(cfg, S, device=local)
    assert pairs.info.fused_available == 1
    stream = torch.cuda.current_stream()
    pairs.set_cuda_stream(stream.cuda_stream)

    # ---- device-resident throughput ("value") ----
    d_audio = torch.from_numpy(audio_np).to(dev)
    d_out = torch.empty((S, F, 2, P, 2), dtype=torch.float32, device=dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()   # nvidia-smi needs ~100 ms to start: launched before the warm-up, read after the timed region
    for _ in range(max(args.warmup, 3)):
        pairs.process_frames(d_audio, hop, F, out=d_out)
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    ev[0].record(stream)
    for i in range(args.steps):
        pairs.process_frames(d_audio, hop, F, out=d_out)
        ev[i + 1].record(stream)
    barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    checksum = float(d_out[:, -1, 0, :, 0].float().mean().item())

    # ---- end to end through the C ABI with HOST buffers ("e2e") ----
    h_audio = torch.from_numpy(audio_np).pin_memory()
    h_out = torch.empty((S, F, 2, P, 2), dtype=torch.float32).pin_memory()
    e2e_steps = 0 if args.no_e2e else max(1, min(args.steps, 3))
    if e2e_steps:
        pairs.process_frames(h_audio.numpy(), hop, F, out=h_out.numpy())  # warm (allocates staging once)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(e2e_steps):
        pairs.process_frames(h_audio.numpy(), hop, F, out=h_out.numpy())
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None   # sampled across the device-resident and the host-buffer timed regions
    e2e_ms = e0.elapsed_time(e1) / max(e2e_steps, 1)

    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0].item()), float(t[1].item())

    if rank == 0:
        frames_per_step = S * F * world
        ms_per_step = total_ms / args.steps
        value = frames_per_step / (ms_per_step * 1e-3)
        peak, peak_src = measured_peak()
        kernel_ms = statistics.mean(step_ms)           # rank 0's own launches
        achieved = S * F * ALGO_BYTES_PER_FRAME / (kernel_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(), "peak_source": peak_src,
                         "algorithmic_bytes_per_frame": ALGO_BYTES_PER_FRAME,
                         "unique_bytes_per_frame": UNIQUE_BYTES_PER_FRAME,
                         "unique_bytes_frac": S * F * UNIQUE_BYTES_PER_FRAME / (kernel_ms * 1e-3) / 1e9 / peak,
                         "kernel": "sgz::k_fused8<8,true,2,0>", "kernel_ms": kernel_ms},
            "e2e": {"value": frames_per_step / (e2e_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(h_audio.numel() * 4), "d2h_bytes_per_step": int(h_out.numel() * 4),
                    "ms_per_step": e2e_ms, "steps": e2e_steps},
            "gpu_launches": args.steps,  # one k_fused8 launch per device-resident step (plus one D2D memcpy node)
            "clocks": clocks, "checksum": checksum,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import binding
            binding.build()
            cores = binding.hardware_concurrency() or os.cpu_count() or 1
            cs, cf = min(max(16 * cores, 64), 1024), args.frames   # ~10-30 s of single-core work
            v, dt = cpu_sample(args, cores, cs, cf)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{cs} streams x {cf} frames of the same C3 workload in {dt:.2f} s, {cores} threads over streams",
                                    "note": "restated reference algorithm (scalar Stockham FFT), not the shipped binary (pffft); parity unpinned"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bind_to_gpu_numa_node(index: int) -> None:
    """Pin this rank's threads to the CPUs next to its GPU (NVML's ideal affinity) before any host buffer is
    allocated, so the pinned staging buffers of the e2e leg are first-touched on the GPU's own NUMA node; with 8
    ranks on a two-socket host the default placement puts half of them across the socket link."""
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(index))
    except Exception:
        pass   # affinity is an optimisation of the host side only


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams", type=int, default=1024)
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--hop", type=int, default=2048)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="kernel A/B runs only: skip the host-buffer leg (the line then carries no e2e number)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
