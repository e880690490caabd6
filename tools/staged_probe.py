"""Throughput of process_frames on configurations that take the staged (four kernels per frame) path."""
import torch, numpy as np
from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch
def run(name, **kw):
    cfg = _abi.default_config(**kw)
    S, F = 1024, 16
    hop = cfg.window_size // 4
    n = signals.frames_length(cfg.window_size, hop, F)
    base = signals.batch(16, n)
    d = torch.from_numpy(np.tile(base, (S // 16, 1, 1))).cuda()
    p = TransformPairBatch(cfg, S); p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
    out = torch.empty((S, F, 2, cfg.axis_points, 2), dtype=torch.float32, device="cuda")
    for _ in range(2): p.process_frames(d, hop, F, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): p.process_frames(d, hop, F, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"{name}: fused={p.info.fused_available} {S * F / ms / 1e3:.2f} M frames/s ({ms:.2f} ms for {S}x{F})")
run("8192 Separate (fused)")
run("8192 Phase", configuration=_abi.CH_PHASE)
run("8192 Complex", configuration=_abi.CH_COMPLEX)
run("16384 Separate", window_size=16384)
run("1024 Separate", window_size=1024)
