"""Independent float64 numpy model of the Spectrum FFT path (mono + Separate/MidSide).

TEST INFRASTRUCTURE ONLY.  A second, separately written restatement of
Source/Spectrum/TransformDSP.inl used to (a) catch transcription errors in the C++ oracle and
(b) budget float32 error: values are float64 (numpy.fft), index decisions replay the reference's
float32 arithmetic (TransformDSP.inl:522-524, :571-577, :606-612).  PARITY UNPINNED like the
C++ oracle (the reference has no tests or golden vectors; cpl is absent)."""
import numpy as np

from signalizer_b200 import _abi

F32 = np.float32

_TERMS = {
    _abi.WIN_RECTANGULAR: [1.0],
    _abi.WIN_HANN: [0.5, 0.5],
    _abi.WIN_HAMMING: [0.54, 0.46],
    _abi.WIN_BLACKMAN: [0.42, 0.5, 0.08],
    _abi.WIN_BLACKMAN_HARRIS: [0.35875, 0.48829, 0.14128, 0.01168],
    _abi.WIN_NUTTALL: [0.355768, 0.487396, 0.144232, 0.012604],
    _abi.WIN_FLATTOP: [0.21557895, 0.41663158, 0.277263158, 0.083578947, 0.006947368],
}


def transform_size(window_size: int) -> int:
    n = 1
    while n < window_size:
        n <<= 1
    return max(32, n)  # TransformConstant.h:84


def window(cfg):
    """-> (float32 kernel[W], scale float32).  generateWindow decision: scale = W / sum(w)."""
    W = cfg.window_size
    D = float(W) if cfg.window_shape == _abi.WINSHAPE_PERIODIC else float(max(W - 1, 1))
    n = np.arange(W, dtype=np.float64)
    w = np.zeros(W)
    for k, a in enumerate(_TERMS[cfg.window_type]):
        w += (-1.0) ** k * a * np.cos(2.0 * np.pi * k * n / D)
    w32 = w.astype(F32)
    return w32, F32(W / w32.astype(np.float64).sum())


def mapped_frequencies(cfg):
    """TransformConstant.h:125-180, non-Complex."""
    P = cfg.axis_points
    i = np.arange(P, dtype=np.float64)
    size = cfg.view_right - cfg.view_left
    fs = float(F32(cfg.sample_rate))
    if cfg.view_scaling == _abi.VIEW_LINEAR:
        half = fs * 0.5
        per = half / (P - 1)
        f = 1.0 * cfg.view_left * half + 1.0 * size * i * per
    else:
        end = fs / 2
        f = cfg.min_freq * np.power(end / cfg.min_freq, cfg.view_left + size * (i / float(P - 1)))
    return f.astype(F32)


def slope_map(cfg, f32):
    a, b = F32(cfg.slope_a), F32(cfg.slope_b)
    return (b * np.power(f32, a).astype(F32)).astype(F32)


def bin_map(cfg):
    """(mode, first, count, fallback) per display point, float32 index arithmetic."""
    N = transform_size(cfg.window_size)
    P = cfg.axis_points
    num_bins = N >> 1
    top = F32(cfg.sample_rate) / F32(2)
    f2b = F32(num_bins / top)  # float32(numBins / topFrequency): int / float -> float
    mf = mapped_frequencies(cfg)
    fft_bw = 1.0 / num_bins
    mode = np.zeros(P, np.int32); first = np.zeros(P, np.int32)
    count = np.zeros(P, np.int32); fallback = np.zeros(P, np.int32)
    x = 0
    while x < P - 1:
        bw = float(F32(F32(mf[x + 1] - mf[x]) / top))
        if bw > fft_bw:
            break
        pos = F32(mf[x] * f2b)
        if cfg.bin_polation == _abi.BIN_NONE:
            first[x] = min(max(int(float(pos) + 0.5), 0), num_bins)
        else:
            first[x] = int(np.floor(pos))
        x += 1
    old = int(F32(mf[x] * f2b))
    while x < P:
        b = int(F32(mf[x] * f2b))
        diff = b - old
        counter = 1 if diff else 0
        mode[x] = 1; first[x] = old + counter; count[x] = max(diff, 1); fallback[x] = b
        old = b
        x += 1
    return mode, first, count, fallback


def _lanczos(v, x32, a):
    """sum_{i=floor(x)-a+1}^{floor(x)+a} v[i] sinc(x-i) sinc((x-i)/a); out-of-range taps 0."""
    x = float(x32)
    base = int(np.floor(x))
    acc = 0.0
    for i in range(base - a + 1, base + a + 1):
        if 0 <= i < len(v):
            d = x - i
            if -a < d < a:
                acc += v[i] * np.sinc(d) * np.sinc(d / a)
    return acc


def _linear(v, x32):
    x = float(x32)
    i = int(np.floor(x)); fr = x - i
    lo = v[i] if 0 <= i < len(v) else 0.0
    hi = v[i + 1] if 0 <= i + 1 < len(v) else 0.0
    return lo * (1 - fr) + hi * fr


def prepare(cfg, audio):
    """audio [n, 2] float32 -> complex128 buffer[N] (TransformDSP.inl:59-223); float32 products."""
    N = transform_size(cfg.window_size)
    W = cfg.window_size
    w, _ = window(cfg)
    L = audio[-W:, 0].astype(F32); R = audio[-W:, 1].astype(F32)
    c = cfg.configuration
    half = F32(0.5)
    if c == _abi.CH_LEFT:
        re, im = L * w, np.zeros(W, F32)
    elif c == _abi.CH_RIGHT:
        re, im = R * w, np.zeros(W, F32)
    elif c == _abi.CH_MERGE:
        re, im = (L + R) * w * half, np.zeros(W, F32)
    elif c == _abi.CH_SIDE:
        re, im = (L - R) * w * half, np.zeros(W, F32)
    elif c == _abi.CH_MIDSIDE:
        re, im = (L + R) * w * half, (L - R) * w * half
    else:
        re, im = L * w, R * w
    buf = np.zeros(N, np.complex128)
    buf[:W] = re.astype(np.float64) + 1j * im.astype(np.float64)
    return buf


def map_to_linear_space(cfg, spectrum):
    """spectrum: complex128[N] (unnormalised DFT of the prepared buffer) -> magnitudes
    [channels, P] float64 (the real parts the reference stores in csp; the imaginary parts are 0
    everywhere except possibly mono bin N/2)."""
    N = transform_size(cfg.window_size)
    P = cfg.axis_points
    W = cfg.window_size
    num_bins = N >> 1
    _, scale = window(cfg)
    inv_size = float(F32(float(scale) / (W * 0.5)))
    top = F32(cfg.sample_rate) / F32(2)
    f2b = F32(num_bins / top)
    mf = mapped_frequencies(cfg)
    mode, first, count, fallback = bin_map(cfg)
    stereo = cfg.configuration in (_abi.CH_SEPARATE, _abi.CH_MIDSIDE)
    Z = np.array(spectrum, dtype=np.complex128)
    if not stereo:
        csf = np.zeros(N + 1, np.complex128)
        csf[:N] = Z
        csf[0] *= 0.5; csf[num_bins] *= 0.5
        csf[:num_bins] = np.abs(csf[:num_bins])
        chans = [csf]
        mirror = [False]
    else:
        k = np.arange(1, num_bins)
        A = 0.5 * (Z[k] + np.conj(Z[N - k]))
        B = (Z[k] - np.conj(Z[N - k])) / 2j
        csf = np.zeros(N + 1, np.complex128)
        csf[k] = A; csf[N - k] = B
        csf[N] = Z[0].imag * 0.5; csf[0] = Z[0].real * 0.5
        csf[num_bins] = Z[num_bins] * 0.5
        csf[num_bins - 1] *= 0.5
        csf[1:N] = np.abs(csf[1:N])
        chans = [csf, csf]
        mirror = [False, True]
    out = np.zeros((len(chans), P))
    for c, (v, mir) in enumerate(zip(chans, mirror)):
        vr = v.real if not np.iscomplexobj(v) else v
        for x in range(P):
            pos = F32(mf[x] * f2b)
            if mode[x] == 0:
                p = F32(F32(N) - pos) if mir else pos
                if cfg.bin_polation == _abi.BIN_NONE:
                    idx = N - first[x] if mir else first[x]
                    val = vr[idx]
                elif cfg.bin_polation == _abi.BIN_LINEAR:
                    val = _linear(vr, p)
                else:
                    val = _lanczos(vr, p, 5)
                out[c, x] = abs(inv_size * val)
            else:
                best, best_bin = 0.0, fallback[x]
                for b in range(first[x], first[x] + count[x]):
                    idx = N - b if mir else b
                    m = abs(vr[idx]) ** 2
                    if m > best:
                        best, best_bin = m, b
                idx = N - best_bin if mir else best_bin
                out[c, x] = abs(inv_size * vr[idx])
    return out


def post_process(cfg, mags, states):
    """mags [channels, P]; states [2 graphs, channels, P] updated in place -> results same shape
    (TransformDSP.inl:1299-1392)."""
    low = 10.0 ** (cfg.low_dbs / 20.0); high = 10.0 ** (cfg.high_dbs / 20.0)
    dy = 1.0 / np.log(high / low)
    mf = mapped_frequencies(cfg)
    slope = slope_map(cfg, mf).astype(np.float64)
    res = np.zeros_like(states)
    for k in range(2):
        pole = float(F32(cfg.pole[k]))
        states[k] *= pole
        np.maximum(states[k], mags, out=states[k])
        d = slope[None, :] * states[k] / low
        with np.errstate(divide="ignore", invalid="ignore"):
            res[k] = np.where(d > 0, np.log(np.where(d > 0, d, 1.0)) * dy, cfg.clip_db)
    return res


def frame(cfg, audio, states):
    """One full frame in float64 -> (mags [ch, P], results [2, ch, P])."""
    buf = prepare(cfg, audio)
    spec = np.fft.fft(buf)
    mags = map_to_linear_space(cfg, spec)
    return mags, post_process(cfg, mags, states)
