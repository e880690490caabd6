"""GPU vs oracle on the Phase configuration at W = 12000 (N = 16384): prints the display points that differ."""
import numpy as np
from oracle.binding import OracleSpectrum
from signalizer_b200 import _abi, signals
from signalizer_b200.spectrum import TransformPairBatch
cfg = _abi.default_config(window_size=12000, window_type=_abi.WIN_HAMMING, axis_points=512, configuration=_abi.CH_PHASE, bin_polation=_abi.BIN_NONE)
S = 3; audio = signals.batch(S, 12000 + 11)
o = OracleSpectrum(cfg, S); g = TransformPairBatch(cfg, S)
o.prepare(audio); o.transform(); ro = o.raw_fft().copy(); o.map()
g.prepare_transform(audio); g.do_transform(); rg = g.get_raw_fft(); g.map_to_linear_space()
to = np.asarray(o.transform_result()).reshape(S, 1024, 2)[:, :512, 0]; tg = g.get_transform_result().reshape(S, 1024, 2)[:, :512, 0]
rel = np.abs(tg - to) / np.maximum(to, 1e-12)
bad = np.argwhere(rel > 1e-3)
print("points off by > 1e-3:", bad.tolist())
mode, first, count, fb = g.bin_map()
for s, x in bad[:4]:
    print("stream", s, "point", x, "mode", mode[x], "first", first[x], "count", count[x], "gpu", tg[s, x], "oracle", to[s, x])
    lo, hi = first[x], first[x] + count[x]
    N = 16384
    zo = ro[s].reshape(-1, 2); zg = rg[s].reshape(-1, 2)
    for k in range(lo, hi):
        a = zo[k, 0] + 1j * zo[k, 1]; b = zo[N - k, 0] - 1j * zo[N - k, 1] if k else 0
        print("   bin", k, "|Z[k]| oracle", abs(zo[k, 0] + 1j * zo[k, 1]), "gpu", abs(zg[k, 0] + 1j * zg[k, 1]), "|Z[N-k]| oracle", abs(zo[N - k, 0] + 1j * zo[N - k, 1]), "gpu", abs(zg[N - k, 0] + 1j * zg[N - k, 1]))
