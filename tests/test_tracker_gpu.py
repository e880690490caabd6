"""Frequency tracker through the C ABI (sgz_spectrum_track_peaks) against the oracle restatement of
Spectrum::drawFrequencyTracking (SpectrumRendering.cpp:301-469) searching the SAME arrays (read back
from the GPU handle), so the integer peak position must agree exactly and the derived doubles to
rounding."""
import numpy as np
import pytest

from oracle import binding as ob
from signalizer_b200 import _abi, _lib, signals
from signalizer_b200.spectrum import TransformPairBatch

pytestmark = pytest.mark.gpu
FIELDS = ["peak_frequency", "peak_dbs", "peak_deviance", "peak_slope", "peak_x", "peak_fraction", "smoothed_frequency", "smoothed_dbs"]


def tones(S, n, fs=48000.0):
    t = np.arange(n) / fs
    a = np.zeros((S, n, 2), np.float32)
    for s in range(S):
        f1, f2 = 220.0 * 1.37 ** s, 3100.3 + 811.0 * s
        a[s, :, 0] = 0.4 * np.sin(2 * np.pi * f1 * t) + 0.1 * np.sin(2 * np.pi * f2 * t)
        a[s, :, 1] = 0.3 * np.sin(2 * np.pi * f2 * t + 0.5)
    return a + 1e-4 * signals.batch(S, n)


def run(cfg, S):
    g = TransformPairBatch(cfg, S)
    audio = tones(S, cfg.window_size)
    if cfg.algo == _abi.ALGO_RSNT:
        g.resonating_dispatch(audio)
    else:
        assert g.prepare_transform(audio)
        g.do_transform()
    g.map_to_linear_space(); g.post_process_std_transform()
    return g


def compare(got, want):
    for s, (a, b) in enumerate(zip(got, want)):
        assert (a.peak_offset, a.searched_raw_fft, a.peak_is_complex) == (b.peak_offset, b.searched_raw_fft, b.peak_is_complex), s
        for f in FIELDS:
            x, y = getattr(a, f), getattr(b, f)
            assert x == y or abs(x - y) <= 1e-9 * max(1.0, abs(y)), (s, f, x, y)


MICE = [0.0, 0.004, 0.2, 0.37, 0.5, 0.61, 0.8, 0.97, 1.0]


@pytest.mark.parametrize("graph", [_abi.TRACK_LINE_MAIN, _abi.TRACK_LINE_SECOND, _abi.TRACK_TRANSFORM])
@pytest.mark.parametrize("kw", [dict(configuration=_abi.CH_LEFT), dict(configuration=_abi.CH_SEPARATE, bin_polation=_abi.BIN_LINEAR),
                                dict(algo=_abi.ALGO_RSNT, window_type=_abi.WIN_HANN), dict(configuration=_abi.CH_MERGE, view_scaling=_abi.VIEW_LINEAR, window_size=4096)])
def test_tracker_matches_oracle_on_the_same_arrays(graph, kw):
    cfg = _abi.default_config(**kw)
    S = len(MICE)
    g = run(cfg, S)
    raw = cfg.algo == _abi.ALGO_FFT and graph == _abi.TRACK_TRANSFORM
    if raw and cfg.configuration == _abi.CH_SEPARATE:
        pytest.skip("audioMemory[N] (Im Z[0] / 2 after the stereo split) is not readable through get_raw_fft")
    got = g.track_peaks(graph, MICE)
    f, sl = g.mapped_frequencies(), g.slope_map()
    if raw:
        rf = np.concatenate([g.get_raw_fft(), np.zeros((S, 1, 2), np.float32)], axis=1)   # element N: the zero pad of prepareTransform
        want = ob.track_peaks(cfg, graph, g.N, g.info.window_kernel_scale, f, sl, MICE, raw_fft=rf)
    else:
        shown = _abi.TRACK_LINE_MAIN if graph == _abi.TRACK_TRANSFORM else graph
        want = ob.track_peaks(cfg, graph, g.N, g.info.window_kernel_scale, f, sl, MICE, results=g.get_results(shown))
    assert all(r.searched_raw_fft == int(raw) for r in got)
    compare(got, want)


def test_raw_tracker_finds_the_tone_and_walks_in_from_a_window_beside_it():
    cfg = _abi.default_config(configuration=_abi.CH_LEFT)
    S = 4
    g = run(cfg, S)
    f = g.mapped_frequencies()
    for s, tone, amp in [(0, 220.0, 0.4), (1, 220.0 * 1.37, 0.4), (2, 220.0 * 1.37 ** 2, 0.4)]:
        x = int(np.argmin(np.abs(f - tone)))
        for dx in (0, 25, -25):                       # mouse on the tone / 25 points beside it (outside +-3 % = 31 points? no: inside)
            mouse = np.full(S, (x + dx) / (cfg.axis_points - 1))
            r = g.track_peaks(_abi.TRACK_TRANSFORM, mouse)[s]
            assert abs(r.peak_frequency - tone) < 0.05 * 48000.0 / 8192, (s, dx, r.peak_frequency)
            assert abs(r.peak_dbs - 20 * np.log10(r.peak_slope) - 20 * np.log10(amp)) < 0.05
    # a window that ends on the rising skirt of the tone: the border walk climbs to the summit
    x = int(np.argmin(np.abs(f - 220.0)))
    r = g.track_peaks(_abi.TRACK_LINE_MAIN, np.full(S, (x - 33) / (cfg.axis_points - 1)))[0]
    res = g.get_results(0)[0, :, 0]
    assert res[r.peak_offset] == res[x - 70: x + 70].max()


def test_tracker_smoothing_state_matches_oracle_over_calls():
    cfg = _abi.default_config(configuration=_abi.CH_LEFT)
    S = 3
    g = run(cfg, S)
    sg, so = (_abi.TrackerState * S)(), (_abi.TrackerState * S)()
    f, sl = g.mapped_frequencies(), g.slope_map()
    res = g.get_results(0)
    for mouse in (0.3, 0.31, 0.6, 0.6, 0.1):
        got = g.track_peaks(_abi.TRACK_LINE_MAIN, mouse, smoothing_ms=120.0, fps=60.0, states=sg)
        want = ob.track_peaks(cfg, 0, g.N, g.info.window_kernel_scale, f, sl, mouse, results=res, smoothing_ms=120.0, fps=60.0, states=so)
        compare(got, want)
    assert bytes(sg) == bytes(so)


def test_tracker_errors():
    cfg = _abi.default_config(configuration=_abi.CH_LEFT)
    g = TransformPairBatch(cfg, 2)
    with pytest.raises(_lib.SgzError) as e:
        g.track_peaks(_abi.TRACK_TRANSFORM, 0.5)      # raw FFT search before any transform
    assert e.value.code == _abi.ESTATE
    with pytest.raises(_lib.SgzError):
        g.track_peaks(7, 0.5)
    g.track_peaks(_abi.TRACK_LINE_MAIN, 0.5)           # results graph: always available (zeros before the first frame)
