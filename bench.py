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


def ncu_traffic(kernel_name):
    """dram bytes per launch of the fused kernel from the committed ncu summary - only when that capture is of the
    kernel this run launches (the summary records the name), otherwise None: a stale profile must not be quoted."""
    path = os.path.join(ROOT, "profiles", "fused_kernel_summary.json")
    try:
        with open(path) as f:
            d = json.load(f)
        captured = d.get("metrics", {}).get("Kernel Name", {}).get("value", "")
        same = kernel_name.replace("sgz::", "").replace(" ", "") in captured.replace("void ", "").replace(" ", "")
        return (d.get("dram_bytes_per_launch"), d.get("capture")) if same and d.get("round") == 2 else (None, None)
    except Exception:
        return None, None


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


def cpu_arm():
    """Which CPU implementation is timed: oracle/_ref (the reference's own TransformDSP.inl / TransformPair.h /
    TransformConstant.h compiled against stand-ins for its absent cpl library) when it was built, else the restated port."""
    from oracle import binding
    binding.build()
    ref = binding.ref_available()
    note = ("the reference's own Source/Spectrum transform code (oracle/_ref), -O3 -mavx2; its cpl primitives (FFT, Lanczos / linear filters, "
            "window) are this repo's stand-ins - a scalar-complex Stockham FFT, not the shipped binary's pffft") if ref else \
           "restated reference algorithm (oracle port, scalar Stockham FFT), not the shipped binary (pffft)"
    return ref, ("reference" if ref else "port"), note


def cpu_sample(args, threads, streams, frames, ref):
    """Time the CPU implementation on a bounded sample of the same workload (threads over streams, like the reference's
    parallel_for over pairs, SpectrumDSP.cpp:83)."""
    from oracle.binding import OracleSpectrum
    from signalizer_b200 import signals
    cfg = bench_cfg()
    n = signals.frames_length(8192, args.hop, frames)
    audio = signals.batch(streams, n)
    o = OracleSpectrum(cfg, streams, fast=True, ref=ref)
    o.process_frames(audio[:, :8192], args.hop, 1, threads=threads, want_results=False)  # touch
    t0 = time.perf_counter()
    o.process_frames(audio, args.hop, frames, threads=threads, want_results=False)
    dt = time.perf_counter() - t0
    return streams * frames / dt, dt


def cpu_stage_split(ref, frames=48):
    """microseconds per frame of the four stages on one core (prepareTransform / doTransform / mapToLinearSpace /
    postProcessStdTransform called one after the other over `frames` single-frame streams)"""
    from oracle.binding import OracleSpectrum
    from signalizer_b200 import signals
    cfg = bench_cfg()
    audio = signals.batch(frames, 8192)
    o = OracleSpectrum(cfg, frames, fast=True, ref=ref)
    best = None
    for _ in range(3):
        t = [time.perf_counter()]
        o.prepare(audio); t.append(time.perf_counter())
        o.transform(); t.append(time.perf_counter())
        o.map(); t.append(time.perf_counter())
        o.postprocess(); t.append(time.perf_counter())
        cur = [(t[i + 1] - t[i]) / frames * 1e6 for i in range(4)]
        best = cur if best is None or sum(cur) < sum(best) else best
    return dict(zip(("prepare_us", "fft_us", "map_us", "post_us"), (round(x, 1) for x in best)))


def cpu_baseline(args):
    from oracle import binding
    ref, kind, note = cpu_arm()
    cores = binding.hardware_concurrency() or os.cpu_count() or 1
    cs, cf = min(max(16 * cores, 64), 1024), args.frames   # ~10-30 s of single-core work
    v, dt = cpu_sample(args, cores, cs, cf, ref)
    v1, dt1 = cpu_sample(args, 1, 16, 16, ref)
    return {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{cs} streams x {cf} frames of the same C3 workload in {dt:.2f} s, {cores} threads over streams",
            "one_core": {"value": v1, "unit": UNIT, "sample": f"16 streams x 16 frames in {dt1:.2f} s on one thread"},
            "stage_split_one_core": cpu_stage_split(ref), "note": note}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path with every host thread.  Its own build system
    cannot run here (External/cpl is absent, JUCE needs X11), so this is oracle/_ref - the reference's Source/Spectrum
    transform sources compiled against cpl stand-ins - or, where that was not built, the restated port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import binding
    ref, kind, note = cpu_arm()
    cores = binding.hardware_concurrency() or os.cpu_count() or 1
    # a step large enough that thread start-up does not count against the reference (about 0.1 s of work on 16 cores),
    # small enough that the default --steps 50 ends within seconds
    streams = max(cores * 16, 64)
    frames = 16
    vals = []
    for i in range(args.warmup + args.steps):
        v, dt = cpu_sample(args, cores, streams, frames, ref)
        if i >= args.warmup:
            vals.append((v, dt))
    value = statistics.median(v for v, _ in vals)
    ms = statistics.median(dt for _, dt in vals) * 1e3
    sample = f"{streams} streams x {frames} frames per step (same C3 config and signals), {cores} threads over streams"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample, "note": note},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from signalizer_b200 import _abi, sharding
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

    def timed(fn, steps):
        """device time of `steps` calls of fn on the launch stream, barrier + synchronize on both sides; per-step list"""
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        barrier()
        ev[0].record(stream)
        for i in range(steps):
            fn()
            ev[i + 1].record(stream)
        barrier()
        return ev[0].elapsed_time(ev[-1]), [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]

    cfg = bench_cfg()
    audio_np, n = make_audio(args, rank)
    S, F, hop, P = args.streams, args.frames, args.hop, cfg.axis_points
    pairs = TransformPairBatch(cfg, S, device=local)
    assert pairs.info.fused_available == 1
    kernel_name = pairs.kernel_name()
    stream = torch.cuda.current_stream()
    pairs.set_cuda_stream(stream.cuda_stream)

    # ---- device-resident throughput ("value"): every rank its own 1024 streams ----
    d_audio = torch.from_numpy(audio_np).to(dev)
    d_out = torch.empty((S, F, 2, P, 2), dtype=torch.float32, device=dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()   # nvidia-smi needs ~100 ms to start: launched before the warm-up, read after the timed region
    for _ in range(max(args.warmup, 3)):
        pairs.process_frames(d_audio, hop, F, out=d_out)
    total_ms, step_ms = timed(lambda: pairs.process_frames(d_audio, hop, F, out=d_out), args.steps)
    checksum = float(d_out[:, -1, 0, :, 0].float().mean().item())

    # ---- the configuration BASELINE.json names at N > 1: 1024 streams in TOTAL, sharded over the ranks ("strong_1024") ----
    lo, hi = sharding.stream_range(rank, world, S)
    strong_ms, gather = total_ms, None
    if world > 1:
        shard = TransformPairBatch(cfg, hi - lo, device=local)
        shard.set_cuda_stream(stream.cuda_stream)
        sh_audio, sh_out = d_audio[lo:hi], d_out[lo:hi]
        for _ in range(max(args.warmup, 3)):
            shard.process_frames(sh_audio, hop, F, out=sh_out)
        strong_ms, _ = timed(lambda: shard.process_frames(sh_audio, hop, F, out=sh_out), args.steps)
        # the optional final gather: one frame-step of results of all 1024 streams to rank 0 over NCCL,
        # outside the metric, reported separately
        last = sh_out[:, -1].contiguous()
        for _ in range(3):
            sharding.gather_results(last, S)
        g_steps = 10
        g_ms, _ = timed(lambda: sharding.gather_results(last, S), g_steps)
        gather = {"bytes": int(S * 2 * P * 2 * 4), "ms": g_ms / g_steps, "what": "last frame of all 1024 streams -> rank 0, torch.distributed gather on NCCL"}

    # ---- end to end through the C ABI with HOST buffers ("e2e" = pinned, full results; the other legs next to it) ----
    e2e_steps = 0 if args.no_e2e else max(1, min(args.steps, 3))
    legs = {}
    if e2e_steps:
        h_audio = torch.from_numpy(audio_np).pin_memory()
        shapes = {"full": ((S, F, 2, P, 2), _abi.OUT_FULL), "line_main": ((S, F, P, 2), _abi.OUT_LINE_MAIN), "last_frame": ((S, 2, P, 2), _abi.OUT_LAST_FRAME)}
        for name, (shape, mode) in shapes.items():
            h_out = torch.empty(shape, dtype=torch.float32).pin_memory()
            call = lambda: pairs.process_frames(h_audio.numpy(), hop, F, out=h_out.numpy(), output=mode)  # noqa: E731
            call()   # warm (sizes the staging buffers once)
            ms, _ = timed(call, e2e_steps)
            legs[name] = (ms / e2e_steps, int(h_audio.numel() * 4), int(h_out.numel() * 4))
        # what the reference's callers hand over: pageable memory (the driver stages it through its own pinned buffers)
        p_out = np.empty((S, F, 2, P, 2), dtype=np.float32)
        call = lambda: pairs.process_frames(audio_np, hop, F, out=p_out)  # noqa: E731
        call()
        ms, _ = timed(call, e2e_steps)
        legs["pageable_full"] = (ms / e2e_steps, int(audio_np.size * 4), int(p_out.size * 4))
    clocks = sampler.stop() if rank == 0 else None   # sampled across the device-resident and the host-buffer timed regions

    keys = sorted(legs)
    t = torch.tensor([total_ms, strong_ms] + [legs[k][0] for k in keys], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, strong_ms = float(t[0].item()), float(t[1].item())
    leg_ms = {k: float(t[2 + i].item()) for i, k in enumerate(keys)}

    if rank == 0:
        frames_per_step = S * F * world
        ms_per_step = total_ms / args.steps
        value = frames_per_step / (ms_per_step * 1e-3)
        peak, peak_src = measured_peak()
        kernel_ms = statistics.mean(step_ms)           # rank 0's own launches
        achieved = S * F * ALGO_BYTES_PER_FRAME / (kernel_ms * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(kernel_name)
        strong_value = S * F / (strong_ms / args.steps * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_frame": ALGO_BYTES_PER_FRAME,
                         "unique_bytes_per_frame": UNIQUE_BYTES_PER_FRAME,
                         "unique_bytes_frac": S * F * UNIQUE_BYTES_PER_FRAME / (kernel_ms * 1e-3) / 1e9 / peak,
                         "kernel": kernel_name, "kernel_ms": kernel_ms},
            # BASELINE.json config 3 as written: 1024 streams in total over the N GPUs (strong scaling; = value at N = 1)
            "strong_1024": {"value": strong_value, "unit": UNIT, "streams_total": S, "streams_per_gpu": hi - lo,
                            "ms_per_step": strong_ms / args.steps,
                            "per_gpu_rate_vs_1024_per_gpu": (strong_value / world) / (value / world), "gather": gather},
            "gpu_launches": args.steps,  # one k_fused8 launch per device-resident step
            "clocks": clocks, "checksum": checksum,
        }
        if legs:
            def leg(k):
                return {"value": frames_per_step / (leg_ms[k] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": legs[k][1],
                        "d2h_bytes_per_step": legs[k][2], "ms_per_step": leg_ms[k], "steps": e2e_steps}
            line["e2e"] = dict(leg("full"), host_memory="pinned", output="full results [S][F][2][P][2]")
            line["e2e_outputs"] = {"line_main": leg("line_main"), "last_frame": leg("last_frame"), "pageable_full": leg("pageable_full")}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
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
