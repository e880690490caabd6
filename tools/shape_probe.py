"""process_frames over (streams, frames) shapes of the C3 geometry on one GPU, device-resident: where the time of a call
goes when there are fewer streams than resident CTAs (296).  Prints us per call, M frames/s, and the excess over the
1024-stream rate.
    PYTHONPATH=. python tools/shape_probe.py [S,F ...]"""
import sys

import numpy as np
import torch

from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch

cfg = _abi.default_config()
hop, W = 2048, 8192
shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]] or [
    (1024, 16), (592, 28), (512, 32), (296, 55), (295, 55), (256, 64), (192, 85), (148, 110), (128, 128), (128, 64), (64, 128), (32, 256), (8, 1024)]
base_rate = None
for S, F in shapes:
    n = signals.frames_length(W, hop, F)
    base = signals.batch(min(S, 16), n)
    reps = (S + base.shape[0] - 1) // base.shape[0]
    d = torch.from_numpy(np.ascontiguousarray(np.tile(base, (reps, 1, 1))[:S])).cuda()
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
    us = e0.elapsed_time(e1) / 20 * 1e3
    rate = S * F / us
    if base_rate is None:
        base_rate = rate
    print(f"{S:5d} streams x {F:4d} frames: {us:8.1f} us  {rate:6.2f} M frames/s  excess over the first shape's rate {us - S * F / base_rate:7.1f} us", flush=True)
    del p, out, d
