"""Resonator bank as a GEMM: the evidence behind DESIGN.md 3.3 (tensor cores for the blocked recurrence).

s[n + B] = p^B s[n] + sum_j p^(B-1-j) x[n + j]: the sum is [2 P V x B] (pole powers, constant) times [B x blocks * channels]
(the samples).  This probe times that contraction with the library GEMM (torch.matmul -> cuBLAS, TF32 tensor cores with a
3-term split, and plain FP32) as a stand-in for the rate a hand-written tcgen05 kernel could reach, checks its accuracy against
float64, and prints the FP32 recurrence kernel's time on the same work (sgz_resonator_process) next to it.

    PYTHONPATH=. python tools/resonator_gemm_probe.py"""
import time

import numpy as np
import torch

from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch


def split_tf32(x):
    hi = x.clone()
    hi.view(torch.int32).bitwise_and_(~0x1FFF)   # keep 10 mantissa bits: exactly representable in TF32
    return hi, x - hi


def main():
    torch.backends.cuda.matmul.allow_tf32 = True
    dev = torch.device("cuda")
    P, n, S = 1024, 480000, 1
    cfg = _abi.default_config(algo=_abi.ALGO_RSNT, axis_points=P)
    g = TransformPairBatch(cfg, S)
    V = g.info.resonator_vectors
    audio = torch.from_numpy(signals.batch(S, n)).to(dev)
    flops = 8.0 * P * V * 2 * n
    states = {}
    for path, name in ((1, "FP32 recurrence (k_resonate_pair + k_resonate_combine)"), (0, "tensor cores (k_resonate_tc + remainder + combine)")):
        g.set_resonator_path(path)
        g.clear_audio_state()
        g.resonating_dispatch(audio)
        states[path] = g.get_resonator_state().astype(np.float64)
        for _ in range(3):
            g.resonating_dispatch(audio)
        g.synchronize()
        t0 = time.perf_counter()
        reps = 10
        for _ in range(reps):
            g.resonating_dispatch(audio)
        g.synchronize()
        ms = (time.perf_counter() - t0) / reps * 1e3
        print(f"{name}: {ms:.3f} ms for {n} samples x {P} filters x {V} vectors x 2 channels = {flops / ms / 1e9:.1f} TFLOP/s-equivalent (8 flop per update)")
    d = np.abs(states[0] - states[1]).max() / np.abs(states[1]).max()
    print(f"state after one pass: max |tensor - fp32| / max |state| = {d:.2e}")

    # pole powers in float64 from random stable poles of the same magnitude range (the rate does not depend on their values)
    rng = np.random.default_rng(1)
    r = 1.0 - 10.0 ** rng.uniform(-4.0, -1.5, size=P * V)
    w = rng.uniform(0.0, np.pi, size=P * V)
    pole = r * np.exp(1j * w)
    for B in (64, 128, 256):
        k = np.arange(B)[::-1]
        pw = pole[:, None] ** k[None, :]                        # [PV, B], p^(B-1-j)
        A64 = np.concatenate([pw.real, pw.imag], axis=0)        # [2PV, B]
        x = audio[0, :n - n % B].T.contiguous()                 # [2, n]
        X64 = x.double().reshape(2 * (x.shape[1] // B), B).T.contiguous()     # [B, blocks*2]
        A, X = torch.from_numpy(A64).float().to(dev), X64.float()
        want = torch.from_numpy(A64).to(dev) @ X64
        scale = float(want.abs().max())
        Ah, Al = split_tf32(A)
        Xh, Xl = split_tf32(X)

        def tf32x3():
            return Ah @ Xh + (Ah @ Xl + Al @ Xh)

        def tf32x1():
            return A @ X

        def fp32():
            torch.backends.cuda.matmul.allow_tf32 = False
            out = A @ X
            torch.backends.cuda.matmul.allow_tf32 = True
            return out

        for name, fn, mults in (("tf32 x3 split", tf32x3, 3), ("tf32 x1", tf32x1, 1), ("fp32 (no tensor cores)", fp32, 1)):
            for _ in range(3):
                out = fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                out = fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            err = float((out.double() - want).abs().max()) / scale
            gf = 2.0 * A.shape[0] * B * X.shape[1]
            print(f"B = {B:3d} {name:24s}: {ms:7.3f} ms  ({gf * mults / ms / 1e9:7.1f} TFLOP/s issued, {flops / ms / 1e9:8.1f} TFLOP/s-equivalent of the recurrence)"
                  f"  max |err| / max |partial| = {err:.2e}   partials {out.numel() * 4 / 1e6:.0f} MB")


if __name__ == "__main__":
    main()
