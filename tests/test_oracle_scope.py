"""Known-answer checks that pin the CPU oracle of the secondary paths (the reference has no tests):
the exact-algebra polar transform against libm, trigger positions of a sine, the Lanczos resampler
on a band-limited tone, the spectral period search on tones."""
import numpy as np

from oracle.binding import OracleScope
from signalizer_b200 import _abi, signals


def test_polar_vertices_equal_libm_atan_sincos():
    o = OracleScope()
    a = signals.batch(3, 5000)
    a[0, :10] = 0
    v = o.vscope_vertices(_abi.vscope_config(gain=1.0), a)
    L, R = a[..., 0].astype(np.float64), a[..., 1].astype(np.float64)
    c, s = -np.sqrt(0.5), np.sqrt(0.5)
    Y, X = L * c - R * s, L * s + R * c
    with np.errstate(all="ignore"):
        ang = np.where((L == 0) & (R == 0), 0.0, np.arctan(X / Y))   # VectorscopeRendering.cpp:580-584
    ln = np.maximum(np.abs(L), np.abs(R))
    assert np.max(np.abs(v[..., 0] - np.sin(ang) * ln)) < 3e-7
    assert np.max(np.abs(v[..., 1] - np.cos(ang) * ln)) < 3e-7
    assert np.all(v[..., 1] >= 0)                                     # atan, not atan2: half-plane fold
    np.testing.assert_allclose(v[0, :, 2], np.arange(5000, dtype=np.float32) * np.float32(1 / 5000) - 1, rtol=0, atol=1e-6)
    # scalePolar: y -> 2 g y - 1, Lissajous: (R, L)
    w = o.vscope_vertices(_abi.vscope_config(scale_polar=True, gain=2.0), a)
    np.testing.assert_allclose(w[..., 1], (v[..., 1] * np.float32(2.0)) * np.float32(2.0) - 1, atol=1e-6)
    z = o.vscope_vertices(_abi.vscope_config(polar=False, gain=1.0), a)
    np.testing.assert_array_equal(z[..., 0], a[..., 1]); np.testing.assert_array_equal(z[..., 1], a[..., 0])


def test_histogram_counts_every_point_inside_the_viewport():
    o = OracleScope()
    a = signals.batch(4, 7000) * np.float32(0.9)
    h = o.vscope_histogram(_abi.vscope_config(gain=1.0, width=64, height=32), a)
    assert h.shape == (4, 32, 64) and np.all(h.sum(axis=(1, 2)) == 7000)
    assert np.all(h[:, :16, :] == 0)   # polar y = cos(.) * len >= 0 -> upper half only
    clipped = o.vscope_histogram(_abi.vscope_config(gain=100.0, width=64, height=32), a)
    assert clipped.sum() < h.sum()


def test_zero_crossing_positions_of_a_sine():
    o = OracleScope()
    n, period = 4800, 96                       # 500 Hz at 48 kHz, phase so that crossings sit between samples
    t = np.arange(n)
    a = np.zeros((1, n, 2), np.float32)
    a[0, :, 0] = np.sin(2 * np.pi * (t - 0.5) / period)
    st = (_abi.ScopeTriggerState * 1)()
    peaks, counts = o.scope_trigger(0, _abi.OSC_LEFT, a, st, threshold=0.5, steady_clock=10)
    # first sample > 0 after a negative one: t = 1 + 96 k (sample 0 is already negative)
    np.testing.assert_array_equal(peaks[0], 10 + 1 + period * np.arange(50))
    assert counts[0] == 50 and st[0].is_peak_holding == 0


def test_lanczos_resampler_reproduces_a_band_limited_tone():
    o = OracleScope()
    n = 3000
    t = np.arange(n)
    a = np.zeros((1, n, 2), np.float32)
    a[0, :, 0] = 0.5 * np.sin(2 * np.pi * 997.0 * t / 48000.0)
    cfg = _abi.resample_config(samples_per_pixel=1 / 16, inc=1 / 4096)
    out = o.scope_resample(cfg, np.array([1000.0]), a)
    m = np.arange(out.shape[1])
    want = 0.5 * np.sin(2 * np.pi * 997.0 * ((n - 1000) + m / 16.0) / 48000.0)
    assert out.shape[1] == 4097 and np.max(np.abs(out[0, :, 1] - want)) < 1e-3
    np.testing.assert_allclose(out[0, ::16, 1], a[0, n - 1000:n - 1000 + 257, 0], atol=1e-6)  # on-sample outputs are the samples


def test_spectral_period_finds_the_fundamental():
    o = OracleScope()
    n = 12000
    t = np.arange(n)
    f0 = np.array([82.4, 440.0, 1000.0])
    a = np.zeros((3, n, 2), np.float32)
    for s in range(3):
        a[s, :, 0] = 0.5 * np.sin(2 * np.pi * f0[s] * t / 48000.0) + 0.3 * np.sin(2 * np.pi * 3 * f0[s] * t / 48000.0)
    prm = _abi.ScopePeriodParams(_abi.ABI_VERSION, _abi.OSC_LEFT, 2048.0, 48000.0, 0.01, 0.05, 0.0)
    st = (_abi.ScopePeriodState * 3)()
    for it in range(6):
        r = o.scope_period(prm, a, st)
        if it < 4:
            assert all(x.record.index == 0 for x in r)   # the zero-initialised 8-deep median wins (OscilloscopeDSP.inl:206-212)
    for s in range(3):
        assert abs(r[s].fundamental - f0[s]) < 0.01 * f0[s]
        assert abs(r[s].cycle_samples - 48000.0 / r[s].fundamental) < 1e-9
        assert -1.0 <= r[s].sample_offset < r[s].cycle_samples


def test_vectorscope_peak_filter_known_answers():
    """VectorScope::runPeakFilter (VectorscopeRendering.cpp:826-889) on constructed inputs"""
    o = OracleScope()
    n = 64
    a = np.zeros((2, n, 2), np.float32)
    a[0, 5, 0], a[0, 9, 1] = -0.5, 0.25          # |L| max 0.5, |R| max 0.25
    a[1, 63, 0] = 0.8                             # stream 1: n = 64 is whole vectors, so the last sample counts
    f = (_abi.VScopeFilters * 2)()
    f, g = o.vscope_peak_envelope(a, f, np.array([1.0, 1.0]), 0.5, 0.25)
    assert tuple(f[0].envelope) == (0.25, 0.0625) and g[0] == 2.0
    assert abs(f[1].envelope[0] - 0.64) < 1e-7 and abs(g[1] - 1.25) < 1e-6
    # second frame, silence: the envelopes decay by coeff = 0.5^(64 * 0.25) = 2^-16, the gain follows
    f, g = o.vscope_peak_envelope(np.zeros_like(a), f, g, 0.5, 0.25)
    assert tuple(f[0].envelope) == (0.25 * 2.0 ** -16, 0.0625 * 2.0 ** -16) and g[0] == 2.0 * 256
    # remainder rule: 65 samples, the loud one is the 65th
    b = np.zeros((1, 65, 2), np.float32)
    b[0, 64, 0] = 1.0
    f1, g1 = o.vscope_peak_envelope(b, (_abi.VScopeFilters * 1)(), np.array([3.0]), 0.5, 0.0)
    assert tuple(f1[0].envelope) == (0.0, 0.0) and g1[0] == 3.0   # 1 / 0 = inf is not normal: the old gain stays
