import time, numpy as np, torch
from signalizer_b200 import _abi, signals
from signalizer_b200.scope import Context
S, n, W = 256, 65536, 4096
d = torch.from_numpy(np.tile(signals.batch(32, n), (8, 1, 1))).cuda()
ctx = Context(); ctx.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
for spp, inc in ((1/16, 1/(16*W)), (1/15, 1/(15*W)), (1/16, 1/(16*W))):
    cfg = _abi.resample_config(channel=_abi.OSC_LEFT, samples_per_pixel=spp, left=0.0, right=1.0, inc=inc)
    t0 = time.perf_counter(); count = ctx.scope_resample_count(cfg); t1 = time.perf_counter()
    pos = np.full(S, W + 100.25)
    out = torch.empty((S, count, 2), dtype=torch.float32, device="cuda")
    for _ in range(3): ctx.scope_resample(cfg, pos, d, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 20
    w0 = time.perf_counter(); e0.record()
    for _ in range(K): ctx.scope_resample(cfg, pos, d, out=out)
    e1.record(); w1 = time.perf_counter(); torch.cuda.synchronize()
    print(f"spp=1/{round(1/spp)} count={count} count_call={1e3*(t1-t0):.3f} ms  host_call={1e3*(w1-w0)/K:.3f} ms  gpu={e0.elapsed_time(e1)/K:.4f} ms per call ({K} calls)")
