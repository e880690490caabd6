"""Per-call latency of process_frames for the real-time shape (few streams, one frame per call), device-resident."""
import time, torch, numpy as np
from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch
for W, wt in ((4096, _abi.WIN_HANN), (8192, _abi.WIN_BLACKMAN_HARRIS)):
    for S in (1, 16, 148):
        cfg = _abi.default_config(window_size=W, window_type=wt)
        d = torch.from_numpy(np.tile(signals.batch(1, W), (S, 1, 1))).cuda()
        p = TransformPairBatch(cfg, S); p.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
        out = torch.empty((S, 1, 2, 1024, 2), dtype=torch.float32, device="cuda")
        for _ in range(20): p.process_frames(d, W, 1, out=out)
        torch.cuda.synchronize()
        K = 200
        t0 = time.perf_counter()
        for _ in range(K): p.process_frames(d, W, 1, out=out)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); p.process_frames(d, W, 1, out=out); e1.record(); torch.cuda.synchronize()
        print(f"window {W} streams {S}: {1e6 * (t1 - t0) / K:.1f} us per call back to back, {1e3 * e0.elapsed_time(e1):.1f} us on the device for one call")
