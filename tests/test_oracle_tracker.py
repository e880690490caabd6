"""Oracle frequency tracker (oracle/sgz_oracle_tracker.hpp) against hand-checkable cases: the search
window, std::max_element's first-maximum rule, the border walk of SpectrumRendering.cpp:320-350
(including its quirk of never stepping onto index 0), and the log-domain parabolic fit of :430-441,
which is exact for a Gaussian peak."""
import numpy as np

from oracle import binding as ob
from signalizer_b200 import _abi


def view(P=1024, **kw):
    cfg = _abi.default_config(axis_points=P, **kw)
    o = ob.OracleSpectrum(cfg, 1)
    return cfg, o.info.transform_size, float(o.info.window_kernel_scale), o.mapped_frequencies().astype(np.float32), o.slope_map().astype(np.float32)


def graph_track(mag, mouse, **kw):
    P = len(mag)
    cfg, N, ws, f, sl = view(P, **kw)
    res = np.zeros((1, P, 2), np.float32); res[0, :, 0] = mag
    return ob.track_peaks(cfg, _abi.TRACK_LINE_MAIN, N, ws, f, sl, mouse, results=res)[0], f, sl, cfg


def test_window_first_maximum_and_fields():
    P = 1000
    mag = np.zeros(P, np.float32); mag[480] = 0.5; mag[495] = 0.75; mag[505] = 0.75; mag[540] = 0.9
    r, f, sl, cfg = graph_track(mag, 0.5)          # pivot 500, range 30 -> [470, 530)
    assert r.peak_offset == 495 and r.searched_raw_fft == 0      # first of the two equal maxima; 540 is outside
    assert r.peak_frequency == float(f[495]) and r.peak_slope == float(sl[495]) and r.peak_x == 495.0
    assert r.peak_dbs == 0.75 * (cfg.high_dbs - cfg.low_dbs) + cfg.low_dbs
    assert r.peak_deviance == float(np.float32(f[496] - f[495]))    # Lanczos: plain neighbour spacing
    r, *_ = graph_track(mag, 0.5, bin_polation=_abi.BIN_LINEAR)
    assert r.peak_deviance == max(float(np.float32(f[496] - f[495])), 0.5 * 8192 / P)


def test_border_walk_climbs_to_the_peak_and_never_reaches_index_zero():
    P = 1000
    x = np.arange(P)
    mag = np.exp(-0.5 * ((x - 600) / 40.0) ** 2).astype(np.float32)   # one broad bump at 600
    assert graph_track(mag, 0.55)[0].peak_offset == 600               # window [520, 580): max on the right border -> walk up
    assert graph_track(mag, 0.66)[0].peak_offset == 600               # window [630, 690): max on the left border -> walk down
    ramp = np.linspace(1.0, 0.0, P).astype(np.float32)                # rising all the way to index 0
    assert graph_track(ramp, 0.10)[0].peak_offset == 1                # the walk stops before results.begin()
    assert graph_track(ramp, 0.01)[0].peak_offset == 0                # lower bound 0: no walk, max_element finds index 0
    up = ramp[::-1].copy()
    assert graph_track(up, 0.9)[0].peak_offset == P - 1               # right walk stops at the last element
    assert graph_track(up, 1.0)[0].peak_deviance > 0                  # offsetIsEnd branch


def test_raw_fft_parabolic_fit_is_exact_for_a_gaussian_peak():
    cfg, N, ws, f, sl = view(1024)
    centre, width, amp = 700.37, 3.0, 0.25
    k = np.arange(N + 1)
    inv = np.float32(ws / (cfg.window_size * 0.5))
    mag = (amp * np.exp(-0.5 * ((k - centre) / width) ** 2)).astype(np.float64) / inv + 1e-30
    raw = np.zeros((1, N + 1, 2), np.float32)
    raw[0, :, 0] = (mag * np.cos(0.3 * k)).astype(np.float32); raw[0, :, 1] = (mag * np.sin(0.3 * k)).astype(np.float32)
    fs = float(cfg.sample_rate)
    target = centre / N * fs
    # mouse fraction whose display frequency is the target (log axis): search there
    mouse = float(np.argmin(np.abs(f - target))) / (cfg.axis_points - 1)
    r = ob.track_peaks(cfg, _abi.TRACK_TRANSFORM, N, ws, f, sl, mouse, raw_fft=raw)[0]
    assert r.searched_raw_fft == 1 and r.peak_offset == 700
    assert abs(r.peak_frequency - target) < 2e-4 * fs / N             # sub-bin position recovered
    slope_db = 20 * np.log10(r.peak_slope)
    assert abs((r.peak_dbs - slope_db) - 20 * np.log10(amp)) < 1e-3
    assert abs(r.peak_fraction - 2 * centre / N) < 1e-6
    want_x = (np.log(target / cfg.min_freq) / np.log(fs / 2 / cfg.min_freq) - cfg.view_left) / (cfg.view_right - cfg.view_left) * (cfg.axis_points - 1)
    assert abs(r.peak_x - want_x) < 1e-3


def test_smoothed_peak_state_converges_and_holds_the_loudest_peak():
    P = 1000
    st = (_abi.TrackerState * 1)()
    cfg, N, ws, f, sl = view(P)
    loud = np.zeros((1, P, 2), np.float32); loud[0, 300, 0] = 0.9
    quiet = np.zeros((1, P, 2), np.float32); quiet[0, 310, 0] = 0.5
    for _ in range(400):
        r = ob.track_peaks(cfg, 0, N, ws, f, sl, 0.3, results=loud, smoothing_ms=50.0, fps=60.0, states=st)[0]
    assert abs(r.smoothed_frequency - float(f[300])) < 1e-6 * float(f[300]) and abs(r.smoothed_dbs - r.peak_dbs) < 1e-6
    r = ob.track_peaks(cfg, 0, N, ws, f, sl, 0.3, results=quiet, smoothing_ms=50.0, fps=60.0, states=st)[0]
    assert r.peak_offset == 310 and st[0].current_frequency == float(f[300])   # the decaying hold still remembers the loud peak
