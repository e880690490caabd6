#!/usr/bin/env python
"""Device-time measurements of the other BASELINE.json configurations (not the driver's contract;
bench.py is).  One JSON line per configuration: C1 (4096-pt Hann frames), C2 (resonator bank,
fp32 FLOP/s), C4 (16x Lanczos resampling + zero-crossing trigger, 256 streams), C5 (polar transform
+ 512x512 histogram, 256 streams).  CUDA events on the launch stream, inputs resident in HBM."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from signalizer_b200 import _abi, signals  # noqa: E402
from signalizer_b200.scope import Context  # noqa: E402
from signalizer_b200.spectrum import TransformPairBatch  # noqa: E402

PEAK = 6551.4
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(fn, sync, steps=10, warmup=3):
    if os.environ.get("SGZ_BENCH_ONCE"):   # one launch per kernel: what an ncu --set full capture of this script wants
        steps, warmup = 1, 0
    for _ in range(warmup):
        fn()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    sync()
    return e0.elapsed_time(e1) / steps


def tile(base, S):
    reps = -(-S // base.shape[0])
    return np.ascontiguousarray(np.tile(base, (reps, 1, 1))[:S])


def c1():
    out_lines = []
    for W in (4096, 2048):
        cfg = _abi.default_config(window_size=W, window_type=_abi.WIN_HANN)
        S, F, hop = 1024, 64, W // 4
        n = signals.frames_length(W, hop, F)
        d = torch.from_numpy(tile(signals.batch(48, n), S)).cuda()
        p = TransformPairBatch(cfg, S)
        p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        out = torch.empty((S, F, 2, 1024, 2), dtype=torch.float32, device="cuda")
        ms = timed(lambda: p.process_frames(d, hop, F, out=out), torch.cuda.synchronize)
        frames = S * F
        bytes_per_frame = 2 * W * 4 + 2 * 2 * 1024 * 8 + 2 * 1024 * 8  # W = 4096: 81 920 B, SURVEY.md §8(d)
        name = "C1 batched" if W == 4096 else "C1 at half the window"
        out_lines.append({"config": f"{name}: 1024 streams x {W}-pt Hann, hop {hop}, Separate, Lanczos, 1024 log bins", "frames_per_s": frames / ms * 1e3,
                          "ms": ms, "achieved_GBps": frames * bytes_per_frame / ms / 1e6, "frac_hbm": frames * bytes_per_frame / ms / 1e6 / PEAK})
    return out_lines


def c3_variants():
    """BASELINE config 3's geometry with the two configurations that have their own fused epilogues."""
    out_lines = []
    for name, conf in (("Phase", _abi.CH_PHASE), ("Complex", _abi.CH_COMPLEX)):
        cfg = _abi.default_config(configuration=conf)
        S, F, hop, W = 1024, 64, 2048, 8192
        n = signals.frames_length(W, hop, F)
        d = torch.from_numpy(tile(signals.batch(48, n), S)).cuda()
        p = TransformPairBatch(cfg, S)
        p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        out = torch.empty((S, F, 2, 1024, 2), dtype=torch.float32, device="cuda")
        ms = timed(lambda: p.process_frames(d, hop, F, out=out), torch.cuda.synchronize)
        frames = S * F
        out_lines.append({"config": f"C3 geometry, {name} configuration: 1024 streams x 8192-pt Blackman-Harris, hop 2048, Lanczos, 1024 log bins",
                          "frames_per_s": frames / ms * 1e3, "ms": ms, "achieved_GBps": frames * 114688 / ms / 1e6, "frac_hbm": frames * 114688 / ms / 1e6 / PEAK})
    return out_lines


def c2():
    cfg = _abi.default_config(algo=_abi.ALGO_RSNT, window_size=8192, window_type=_abi.WIN_HANN)
    out = []
    for S, n in ((1, 480000), (64, 48000)):
        d = torch.from_numpy(tile(signals.batch(min(S, 16), n), S)).cuda()
        p = TransformPairBatch(cfg, S)
        p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        V = p.info.resonator_vectors
        flops = 8.0 * 1024 * V * 2 * n * S   # 8 flop per complex one-pole update, filters x vectors x channels
        for path, name in ((0, "tensor cores (k_resonate_tc, TF32 x3 split)"), (1, "FP32 recurrence (k_resonate_pair)")):
            p.set_resonator_path(path)
            ms = timed(lambda: p.resonating_dispatch(d), torch.cuda.synchronize, steps=5)
            out.append({"config": f"C2 resonator bank, {name}: {S} stereo stream(s), 1024 log bins, Hann (V={V}), {n} samples", "ms": ms,
                        "samples_per_s_per_stream": n / ms * 1e3, "realtime_factor_48k": n / ms * 1e3 / 48000.0,
                        "recurrence_equivalent_TFLOPs": flops / ms / 1e9})
    return out


def c4():
    S, n = 256, 65536
    d = torch.from_numpy(tile(signals.batch(32, n), S)).cuda()
    ctx = Context()
    ctx.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
    W = 4096  # window samples shown
    cfg = _abi.resample_config(channel=_abi.OSC_LEFT, samples_per_pixel=1.0 / 16, left=0.0, right=1.0, inc=1.0 / (16 * W))
    count = ctx.scope_resample_count(cfg)
    pos = np.full(S, W + 100.25)
    out = torch.empty((S, count, 2), dtype=torch.float32, device="cuda")
    ms = timed(lambda: ctx.scope_resample(cfg, pos, d, out=out), torch.cuda.synchronize)
    algo = S * (W * 4 + count * 8)   # read W samples of one channel, write 16 W vertices (x, y)
    st = (_abi.ScopeTriggerState * S)()
    ms_t = timed(lambda: ctx.scope_trigger(0, _abi.OSC_LEFT, d, st, 0.05, max_peaks=4096), torch.cuda.synchronize, steps=5)
    # the same trigger with states / peaks / counts resident on the device (sgz_scope_trigger_device): no copies, no wait
    import ctypes as C
    d_states = torch.zeros((S, C.sizeof(_abi.ScopeTriggerState)), dtype=torch.uint8, device="cuda")
    d_peaks = torch.zeros((S, 4096), dtype=torch.int64, device="cuda")
    d_counts = torch.zeros(S, dtype=torch.int32, device="cuda")
    ms_td = timed(lambda: ctx.scope_trigger_device(0, _abi.OSC_LEFT, d, d_states, d_peaks, d_counts, 0.05), torch.cuda.synchronize, steps=10)
    # calculateFundamentalPeriod + calculateTriggeringOffset (a19): fp64 8192-point FFT, peak pick, Goertzel, per stream
    prm = _abi.ScopePeriodParams(_abi.ABI_VERSION, _abi.OSC_LEFT, 2048.0, 48000.0, 0.01, 0.05, 0.0)
    pst = (_abi.ScopePeriodState * S)()
    ms_p = timed(lambda: ctx.scope_period(prm, d, pst), torch.cuda.synchronize, steps=5)
    return {"config": f"C4 oscilloscope: 256 streams, Lanczos-10 16x resampling of a {W}-sample window ({count} vertices) + zero-crossing trigger over {n} samples",
            "resample_ms": ms, "vertices_per_s": S * count / ms * 1e3, "resample_achieved_GBps": algo / ms / 1e6, "resample_frac_hbm": algo / ms / 1e6 / PEAK,
            "trigger_ms_incl_host_copies": ms_t, "trigger_samples_per_s": S * n / ms_t * 1e3,
            "trigger_device_arrays_ms": ms_td, "trigger_device_arrays_samples_per_s": S * n / ms_td * 1e3,
            "period_search_ms_256_streams_incl_host_tail": ms_p, "period_searches_per_s": S / ms_p * 1e3}


def c4_colour():
    out = []
    for S, n in ((256, 65536), (4096, 8192)):
        d = torch.from_numpy(tile(signals.batch(32, n), S)).cuda()
        ctx = Context()
        ctx.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        cfg = _abi.colour_config(channel_mode=_abi.OSC_SEPARATE, envelope_mode=_abi.ENVELOPE_RMS)
        st = (_abi.ScopeColourState * (2 * S))()
        rgba = torch.empty((S, 4, n, 4), dtype=torch.uint8, device="cuda")
        ms = timed(lambda: ctx.scope_colour(cfg, d, st, out=rgba), torch.cuda.synchronize, steps=3, warmup=1)
        out.append({"config": f"C4 oscilloscope colouring: {S} stereo streams x {n} samples, 3-band Linkwitz-Riley + 4 colour planes + RMS envelope",
                    "ms": ms, "samples_per_s": S * n / ms * 1e3, "algorithmic_GBps": S * n * 24 / ms / 1e6, "frac_hbm": S * n * 24 / ms / 1e6 / PEAK,
                    "note": "recursive filters: one dependent chain per channel, parallel over streams only"})
    return out


def c5():
    S, n = 256, 65536
    d = torch.from_numpy(tile(signals.batch(32, n), S)).cuda()
    ctx = Context()
    ctx.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
    cfg = _abi.vscope_config(gain=1.2)
    hist = torch.empty((S, 512, 512), dtype=torch.int32, device="cuda")
    ms = timed(lambda: ctx.vscope_histogram(cfg, d, out=hist), torch.cuda.synchronize)
    algo = S * (2 * n * 4 + 512 * 512 * 4 * 2)   # read audio, clear + write-out the 1 MiB histogram
    return {"config": f"C5 vectorscope: 256 stereo streams x {n} samples, polar transform + 512x512 histogram", "ms": ms,
            "points_per_s": S * n / ms * 1e3, "achieved_GBps": algo / ms / 1e6, "frac_hbm": algo / ms / 1e6 / PEAK}


def c3_few_streams():
    """BASELINE config 3 cut for N GPUs: 1024 / N streams per GPU (the strong-scaling shape) and long recordings; fewer
    streams than the GPU holds CTAs, so parts start from zero and the seam kernels carry the hold across"""
    out_lines = []
    cfg = _abi.default_config()
    for S, F in ((512, 64), (256, 64), (128, 64), (8, 1024), (1, 4096)):
        hop, W = 2048, 8192
        n = signals.frames_length(W, hop, F)
        d = torch.from_numpy(tile(signals.batch(min(S, 16), n), S)).cuda()
        p = TransformPairBatch(cfg, S)
        p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        out = torch.empty((S, F, 2, 1024, 2), dtype=torch.float32, device="cuda")
        ms = timed(lambda: p.process_frames(d, hop, F, out=out), torch.cuda.synchronize)
        out_lines.append({"config": f"C3 with {S} streams x {F} frames on one GPU", "frames_per_s": S * F / ms * 1e3, "ms": ms,
                          "frac_hbm": S * F * 114688 / ms / 1e6 / PEAK})
    return out_lines


if __name__ == "__main__":
    for fn in (c1, c3_variants, c3_few_streams, c2, c4, c4_colour, c5):
        r = fn()
        for line in (r if isinstance(r, list) else [r]):
            print(json.dumps(line), flush=True)
