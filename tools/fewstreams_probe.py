"""Fixed cost of a process_frames call when there are fewer streams than resident CTAs (the per-GPU shape of BASELINE
config 3 on several GPUs): 128 streams x F frames for several F; the slope is the frame rate, the intercept the per-call
overhead (launches, prologue, cold first frames, seam fix-up, the copy of the last frame).
    PYTHONPATH=. python tools/fewstreams_probe.py"""
import numpy as np
import torch

from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch

cfg = _abi.default_config()
S, hop, W = 128, 2048, 8192
rows = []
for F in (4, 8, 16, 32, 64, 128, 256):
    n = signals.frames_length(W, hop, F)
    base = signals.batch(16, n)
    d = torch.from_numpy(np.ascontiguousarray(np.tile(base, (S // 16, 1, 1)))).cuda()
    p = TransformPairBatch(cfg, S)
    p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
    out = torch.empty((S, F, 2, 1024, 2), dtype=torch.float32, device="cuda")
    for _ in range(3):
        p.process_frames(d, hop, F, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        p.process_frames(d, hop, F, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    rows.append((S * F, ms))
    print(f"{S} streams x {F:3d} frames: {ms * 1e3:8.1f} us  ({S * F / ms / 1e3:6.2f} M frames/s)")
x = np.array([r[0] for r in rows], dtype=np.float64); y = np.array([r[1] for r in rows]) * 1e3
slope, icpt = np.polyfit(x, y, 1)
print(f"fit: {icpt:.1f} us per call + {slope * 1e3:.3f} ns per frame ({1e3 / slope / 1e6:.2f} M frames/s asymptotic)")
