"""Oracle restatement of the oscilloscope colouring / envelopes (oracle/sgz_oracle_colour.hpp) against
properties that hold for any correct Linkwitz-Riley network and for the reference's envelope recursions."""
import numpy as np

from oracle.binding import OracleScope
from signalizer_b200 import _abi


def states(S):
    return (_abi.ScopeColourState * (2 * S))()


def tone(freq, n, amp=0.5, fs=48000.0, ch=(1.0, 1.0)):
    t = np.arange(n) / fs
    a = np.zeros((1, n, 2), np.float32)
    a[0, :, 0] = ch[0] * amp * np.sin(2 * np.pi * freq * t)
    a[0, :, 1] = ch[1] * amp * np.sin(2 * np.pi * freq * t)
    return a


def test_band_colours_follow_the_tone_frequency():
    # pure colours per band, no key blending: a 60 Hz tone paints red (low), 1 kHz green (mid), 9 kHz blue (high)
    o = OracleScope()
    for freq, want in ((60.0, 0), (1000.0, 1), (9000.0, 2)):
        cfg = _abi.colour_config(colouring_blend=1.0, colour_smoothing_ms=5.0)
        rgba, _ = o.scope_colour(cfg, tone(freq, 24000), states(1))
        px = rgba[0, :, -1]                       # last sample of the four planes (left, right, mid, side)
        for plane in (0, 1, 2):
            assert px[plane, want] == 255 and px[plane, 3] == 255, (freq, plane, px[plane])
            others = [c for c in range(3) if c != want]
            assert all(px[plane, c] < 8 for c in others), (freq, plane, px[plane])


def test_side_plane_is_silent_for_identical_channels_and_blend_reaches_the_key():
    o = OracleScope()
    cfg = _abi.colour_config(colouring_blend=1.0)
    rgba, _ = o.scope_colour(cfg, tone(1000.0, 4096), states(1))
    assert np.all(rgba[0, 3, :, :3] == 0) and np.all(rgba[0, 3, :, 3] == 255)   # side bands are exactly zero: 0 * inf = NaN -> 0
    cfg = _abi.colour_config(colouring_blend=0.0, key_left=(10, 20, 30, 40), key_right=(50, 60, 70, 80))
    rgba, _ = o.scope_colour(cfg, tone(1000.0, 512), states(1))
    assert np.all(rgba[0, 0] == (10, 20, 30, 40)) and np.all(rgba[0, 2] == (10, 20, 30, 40))   # blend = 1 - 0: the key everywhere
    assert np.all(rgba[0, 1] == (50, 60, 70, 80)) and np.all(rgba[0, 3] == (50, 60, 70, 80))


def crossover(x, fs=48000.0):
    import ctypes as C
    from oracle import binding as ob
    lib = ob.load()
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty((len(x), 3), np.float32)
    lib.sgzo_scope_crossover.restype = None
    lib.sgzo_scope_crossover.argtypes = [C.c_double, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.sgzo_scope_crossover(fs, x.ctypes.data, len(x), out.ctypes.data)
    return out


def test_crossover_matches_scipy_linkwitz_riley_and_sums_to_an_allpass():
    from scipy import signal
    fs = 48000.0
    rng = np.random.default_rng(3)
    x = rng.standard_normal(20000).astype(np.float32)
    got = crossover(x, fs).astype(np.float64)
    lp1, hp1 = signal.butter(2, 300.0, "low", fs=fs, output="sos"), signal.butter(2, 300.0, "high", fs=fs, output="sos")
    lp2, hp2 = signal.butter(2, 3000.0, "low", fs=fs, output="sos"), signal.butter(2, 3000.0, "high", fs=fs, output="sos")
    twice = lambda sos, v: signal.sosfilt(sos, signal.sosfilt(sos, v))
    low, rest = twice(lp1, x.astype(np.float64)), twice(hp1, x.astype(np.float64))
    want = np.stack([twice(lp2, low) + twice(hp2, low), twice(lp2, rest), twice(hp2, rest)], axis=1)
    assert np.max(np.abs(got - want)) < 2e-4 * np.max(np.abs(want))      # float32 recursions vs float64
    # Linkwitz-Riley: the bands add up to an allpass -> the summed impulse response has |H(f)| = 1
    imp = np.zeros(32768, np.float32); imp[0] = 1.0
    H = np.abs(np.fft.rfft(crossover(imp, fs).astype(np.float64).sum(axis=1)))
    assert np.max(np.abs(H - 1.0)) < 2e-3


def test_rms_envelope_and_gain():
    o = OracleScope()
    n = 48000
    a = tone(440.0, n, amp=0.5, ch=(1.0, 0.5))
    for mode, want in ((_abi.OSC_LEFT, (0.125, 0.125)), (_abi.OSC_RIGHT, (0.03125, 0.03125)), (_abi.OSC_MID, (0.0703125, 0.0703125)),
                       (_abi.OSC_SIDE, (0.0078125, 0.0078125)), (_abi.OSC_SEPARATE, (0.125, 0.03125)), (_abi.OSC_MIDSIDE, (0.140625, 0.015625))):
        st = states(1)
        cfg = _abi.colour_config(channel_mode=mode, envelope_mode=_abi.ENVELOPE_RMS, envelope_window=0.05)
        _, gain = o.scope_colour(cfg, a, st)
        env = (st[0].envelope, st[1].envelope)
        assert abs(env[0] - want[0]) < 0.02 * want[0] and abs(env[1] - want[1]) < 0.02 * want[1], (mode, env)
        assert abs(gain[0] - 1.0 / np.sqrt(max(env))) < 1e-6 * gain[0]
    st = states(1)
    cfg = _abi.colour_config(envelope_mode=_abi.ENVELOPE_PEAK)      # not RMS: envelope computed but not stored, gain 0
    _, gain = o.scope_colour(cfg, a, st)
    assert st[0].envelope == 0 and gain[0] == 0


def test_peak_filter():
    o = OracleScope()
    n = 1003                                                          # the last 3 samples are outside the whole 8-vectors
    a = np.zeros((1, n, 2), np.float32)
    a[0, 100, 0], a[0, 200, 1] = -0.5, 0.25
    a[0, 1001, 0] = 0.9                                               # ignored (remainder)
    for mode, want in ((_abi.OSC_LEFT, (0.25, 0.25)), (_abi.OSC_RIGHT, (0.0625, 0.0625)), (_abi.OSC_MID, (0.0625, 0.0625)),
                       (_abi.OSC_SEPARATE, (0.25, 0.25)),             # the reference never resets its running maximum between channels
                       (_abi.OSC_MIDSIDE, (0.0625, 0.0625))):
        env = np.zeros((1, 2), np.float32)
        gain = o.scope_peak_envelope(mode, a, env, 0.9)
        assert tuple(env[0]) == want, (mode, env)
        assert gain[0] == 1.0 / np.sqrt(np.float32(max(want)))
    env = np.array([[4.0, 1.0]], np.float32)                          # a decaying louder past wins over the window maximum
    o.scope_peak_envelope(_abi.OSC_LEFT, a, env, 0.5)
    assert tuple(env[0]) == (2.0, 0.5)
