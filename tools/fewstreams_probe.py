"""Throughput with few streams and many frames (a long recording), device-resident."""
import os, torch, numpy as np
from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch
cfg = _abi.default_config()
for S, F in ((1, 4096), (8, 1024)):
    hop = 2048
    n = signals.frames_length(8192, hop, F)
    d = torch.from_numpy(signals.batch(S, n)).cuda()
    out = torch.empty((S, F, 2, 1024, 2), dtype=torch.float32, device="cuda")
    for env in ("", "1"):
        if env: os.environ["SGZ_NO_FRAME_PARALLEL"] = env
        else: os.environ.pop("SGZ_NO_FRAME_PARALLEL", None)
        p = TransformPairBatch(cfg, S); p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        for _ in range(2): p.process_frames(d, hop, F, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); p.process_frames(d, hop, F, out=out); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"{S} stream(s) x {F} frames, {'one CTA per stream' if env else 'frame-parallel'}: {ms:.3f} ms = {S * F / ms / 1e3:.2f} M frames/s")
