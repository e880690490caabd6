"""Parity of the CUDA path (through the C ABI) against the CPU oracle on seeded inputs.
Run on the B200 box: python -m pytest tests -m gpu"""
import numpy as np
import pytest

from oracle.binding import OracleSpectrum
from signalizer_b200 import _abi, _lib, signals
from signalizer_b200.spectrum import TransformPairBatch
from tolerances import FLOOR, REL, assert_linear_close, assert_log_close

pytestmark = pytest.mark.gpu

ALL_CONFS = list(range(8))
INTERPS = [_abi.BIN_NONE, _abi.BIN_LINEAR, _abi.BIN_LANCZOS]


def complex_close(got, want, what, rel=REL, floor=FLOOR):
    """[..., 2] (re, im) arrays; distance in the complex plane, per-stream peak."""
    g = got[..., 0].astype(np.float64) + 1j * got[..., 1].astype(np.float64)
    w = want[..., 0].astype(np.float64) + 1j * want[..., 1].astype(np.float64)
    peak = np.abs(w).reshape(w.shape[0], -1).max(axis=1).reshape((-1,) + (1,) * (w.ndim - 1))
    bound = rel * np.abs(w) + floor * peak
    err = np.abs(g - w)
    bad = err > bound
    if np.any(bad):
        i = np.unravel_index(np.argmax(err - bound), err.shape)
        raise AssertionError(f"{what}: {bad.sum()} of {bad.size} outside tolerance; worst at {i}: got {g[i]} want {w[i]} err {err[i]:.3e} bound {bound[i]:.3e}")


def run_staged(obj, audio, oracle=False):
    if oracle:
        assert obj.prepare(audio) == _abi.OK
        obj.transform(); raw = obj.raw_fft().copy(); obj.map(); tr = obj.transform_result().copy(); obj.postprocess()
        return raw, tr, [obj.results(0), obj.results(1)], [obj.states(0), obj.states(1)]
    assert obj.prepare_transform(audio)
    obj.do_transform(); raw = obj.get_raw_fft(); obj.map_to_linear_space(); tr = obj.get_transform_result(); obj.post_process_std_transform()
    return raw, tr, [obj.get_results(0), obj.get_results(1)], [obj.get_states(0), obj.get_states(1)]


SIZES = [(4096, _abi.WIN_HANN, 1024), (8192, _abi.WIN_BLACKMAN_HARRIS, 1024), (700, _abi.WIN_BLACKMAN, 300),
         (20, _abi.WIN_HANN, 40), (20000, _abi.WIN_NUTTALL, 512), (400, _abi.WIN_HANN, 128), (10000, _abi.WIN_HANN, 512), (40000, _abi.WIN_BLACKMAN_HARRIS, 700)]
# transform sizes 4096, 8192 (k_fft8), 1024, 32, 512 (shared-memory FFT, even and odd log2), 16384, 32768, 65536 (1, 2, 3
# global passes, then 8192-point sub-transforms in k_fft8)
# (W = 12000 Hamming would also give 16384, but in the Phase configuration stream 1 has two bins, 3238 and 3241, whose
# channel-2 magnitudes tie to 1.1e-6 inside one max-of-bins range: which one wins depends on the FFT's rounding,
# tools/phase_probe.py)


@pytest.mark.parametrize("size", SIZES)
@pytest.mark.parametrize("conf", ALL_CONFS)
@pytest.mark.parametrize("bp", INTERPS)
def test_staged_pipeline_matches_oracle(size, conf, bp):
    W, wt, P = size
    cfg = _abi.default_config(window_size=W, window_type=wt, axis_points=P, configuration=conf, bin_polation=bp)
    S = 3
    audio = signals.batch(S, W + 11)
    o = OracleSpectrum(cfg, S)
    g = TransformPairBatch(cfg, S)
    phase = conf == _abi.CH_PHASE
    for rep in range(2):  # second pass exercises the peak-hold carry
        raw_o, tr_o, res_o, st_o = run_staged(o, audio if rep == 0 else audio * np.float32(0.25), oracle=True)
        raw_g, tr_g, res_g, st_g = run_staged(g, audio if rep == 0 else audio * np.float32(0.25))
        # raw FFT: error relative to the spectrum's largest bin (float32 FFT error model)
        complex_close(raw_g, raw_o, f"raw FFT rep {rep}", rel=2e-6, floor=2e-6)
        ch = o.channels
        tg = tr_g.reshape(S, ch, P, 2); to = tr_o.reshape(S, ch, P, 2)
        if not phase:
            used = 1 if conf in (_abi.CH_LEFT, _abi.CH_RIGHT, _abi.CH_MERGE, _abi.CH_SIDE, _abi.CH_COMPLEX) else 2
            complex_close(tg[:, :used], to[:, :used], f"transform result rep {rep}")
            lin = np.hypot(to[:, :used, :, 0], to[:, :used, :, 1])
        else:
            # wsp[2x] = magnitude sum, wsp[2x+1] = 1 - cancellation ratio
            mag_g, mag_o = tg.reshape(S, 2 * P, 2)[:, :P, 0], to.reshape(S, 2 * P, 2)[:, :P, 0]
            assert_linear_close(mag_g, mag_o, frame_axes=(1,), what="phase magnitudes")
            can_g, can_o = tg.reshape(S, 2 * P, 2)[:, :P, 1], to.reshape(S, 2 * P, 2)[:, :P, 1]
            strong = mag_o > 1e-3 * mag_o.max(axis=1, keepdims=True)
            strong[:, -1] = False  # never-broke edge: the reference leaves the last cancellation stale
            assert np.all(np.abs(can_g[strong] - can_o[strong]) < 2e-3)
            lin = None
        for k in range(2):
            if not phase:
                assert_linear_close(st_g[k][..., :used], st_o[k][..., :used], frame_axes=(1, 2), what=f"states {k} rep {rep}")
                st_lin = st_o[k][..., :used].astype(np.float64)
                mask = st_lin > 0
                peak = st_lin.max()
                assert_log_close(res_g[k][..., :used][mask], res_o[k][..., :used][mask], lin=st_lin[mask], peak=peak, what=f"results {k} rep {rep}")
                zero = ~mask
                assert np.all(res_g[k][..., :used][zero] == res_o[k][..., :used][zero])
            else:
                assert_linear_close(st_g[k][..., 0], st_o[k][..., 0], frame_axes=(1,), what=f"phase states {k}")


def frames_case(cfg, S, F, hop, layout=_abi.LAYOUT_INTERLEAVED, first_stream=0, scale=1.0):
    n = signals.frames_length(cfg.window_size, hop, F)
    audio = signals.batch(S, n, first_stream=first_stream) * np.float32(scale)
    if layout == _abi.LAYOUT_PLANAR:
        audio = np.ascontiguousarray(audio.transpose(0, 2, 1))
    return audio


def compare_frames(got, want, cfg, what, peak_floor=None):
    """results [S, F, 2, P, 2] on the display scale: recover the linear peak-held state
    v = low * exp(y / dy) from the oracle to scale the tolerance per (stream, frame)."""
    dy = 1.0 / np.log(10.0 ** (cfg.high_dbs / 20.0) / 10.0 ** (cfg.low_dbs / 20.0))
    low = 10.0 ** (cfg.low_dbs / 20.0)
    want64 = want.astype(np.float64)
    clipped = want == np.float32(cfg.clip_db)
    lin = np.where(clipped, 0.0, low * np.exp(np.where(clipped, 0.0, want64) / dy))
    S, F = want.shape[:2]
    for s in range(S):
        for f in range(F):
            w = want[s, f]; g = got[s, f]; l = lin[s, f]
            m = ~clipped[s, f]
            assert np.all(g[~m] == w[~m]) or np.all(l[~m] == 0), f"{what}: clip mismatch stream {s} frame {f}"
            if m.any():
                # the float32 FFT's error floor follows the largest bin of the whole spectrum, which a zoomed view may not show
                peak = l.max() if peak_floor is None else max(l.max(), peak_floor[s])
                assert_log_close(g[m], w[m], lin=l[m], peak=peak, dy=dy, what=f"{what} stream {s} frame {f}")


FUSED_SIZES = [(8192, _abi.WIN_BLACKMAN_HARRIS), (4096, _abi.WIN_HANN), (6000, _abi.WIN_HAMMING), (3000, _abi.WIN_BLACKMAN),
               (2048, _abi.WIN_HANN), (1500, _abi.WIN_HAMMING), (1024, _abi.WIN_BLACKMAN_HARRIS), (700, _abi.WIN_HANN),
               (16384, _abi.WIN_BLACKMAN_HARRIS), (10000, _abi.WIN_HANN)]
# N = 8192, 4096, 8192, 4096, 2048, 2048, 1024, 1024, 16384, 16384


@pytest.mark.parametrize("size", FUSED_SIZES)
@pytest.mark.parametrize("conf", [_abi.CH_LEFT, _abi.CH_RIGHT, _abi.CH_MERGE, _abi.CH_SIDE, _abi.CH_SEPARATE, _abi.CH_MIDSIDE, _abi.CH_COMPLEX])
@pytest.mark.parametrize("bp", INTERPS)
def test_fused_frames_match_oracle(size, conf, bp):
    cfg = _abi.default_config(window_size=size[0], window_type=size[1], configuration=conf, bin_polation=bp)
    S, F, hop = 5, 6, size[0] // 4
    g = TransformPairBatch(cfg, S)
    assert g.info.fused_available == 1
    audio = frames_case(cfg, S, F, hop)
    o = OracleSpectrum(cfg, S)
    want = o.process_frames(audio, hop, F)
    got = g.process_frames(audio, hop, F)
    used = 2 if conf in (_abi.CH_SEPARATE, _abi.CH_MIDSIDE) else 1
    compare_frames(got[..., :used], want[..., :used], cfg, "fused")
    if used == 1:
        assert np.all(got[..., 1] == 0)  # results[i].phase = 0 (TransformDSP.inl:1347)
    # states carried out of the call and the handle's last results
    for k in range(2):
        assert_linear_close(g.get_states(k)[..., :used], o.states(k)[..., :used], frame_axes=(1, 2), what=f"final states {k}")
        np.testing.assert_array_equal(g.get_results(k), got[:, -1, k])


@pytest.mark.parametrize("win", range(_abi.WIN_END))
@pytest.mark.parametrize("shape", [_abi.WINSHAPE_SYMMETRIC, _abi.WINSHAPE_PERIODIC])
@pytest.mark.parametrize("rate,window", [(44100.0, 4096), (96000.0, 8192), (48000.0, 2048)])
def test_fused_window_families_and_sample_rates(win, shape, rate, window):
    cfg = _abi.default_config(window_size=window, window_type=win, window_shape=shape, sample_rate=rate, min_freq=20.0)
    cfg.pole[0], cfg.pole[1] = 0.5, 0.999
    S, F, hop = 2, 4, window // 8
    g = TransformPairBatch(cfg, S)
    assert g.info.fused_available == 1
    audio = frames_case(cfg, S, F, hop)
    got = g.process_frames(audio, hop, F)
    want = OracleSpectrum(cfg, S).process_frames(audio, hop, F)
    compare_frames(got, want, cfg, f"window type {win} shape {shape} rate {rate}")


@pytest.mark.parametrize("bp", INTERPS)
@pytest.mark.parametrize("window", [8192, 1024])
@pytest.mark.parametrize("view", [dict(view_scaling=_abi.VIEW_LINEAR), dict(), dict(view_scaling=_abi.VIEW_LINEAR, view_left=0.0, view_right=0.002)])
def test_fused_complex_configuration_keeps_the_dc_bin_complex(bp, window, view):
    # Complex leaves csf[0] a complex number (halved) among magnitudes; with a DC offset on both channels the display
    # points whose taps reach bin 0 depend on its imaginary part
    cfg = _abi.default_config(window_size=window, configuration=_abi.CH_COMPLEX, bin_polation=bp, **view)
    S, F, hop = 3, 2, window // 2
    g = TransformPairBatch(cfg, S)
    assert g.info.fused_available == 1
    audio = frames_case(cfg, S, F, hop) * np.float32(0.2)
    audio[..., 0] += np.float32(0.3); audio[..., 1] -= np.float32(0.2)
    got = g.process_frames(audio, hop, F)
    want = OracleSpectrum(cfg, S).process_frames(audio, hop, F)
    compare_frames(got[..., :1], want[..., :1], cfg, f"complex dc {view} bp {bp}")
    assert np.all(got[..., 1] == 0)


@pytest.mark.parametrize("bp", INTERPS)
@pytest.mark.parametrize("size", [(8192, _abi.WIN_BLACKMAN_HARRIS), (4096, _abi.WIN_HANN), (3000, _abi.WIN_BLACKMAN), (2048, _abi.WIN_HANN), (1024, _abi.WIN_HAMMING), (16384, _abi.WIN_HANN)])
@pytest.mark.parametrize("view", [dict(), dict(view_scaling=_abi.VIEW_LINEAR, view_left=0.0, view_right=0.1)])
def test_fused_phase_configuration(bp, size, view):
    """Phase: magnitude sum (graph .x) against the oracle like every other configuration; the smoothed
    cancellation (.y) against the oracle and the staged GPU path with the looser bound the staged test uses
    (a ratio of two nearly equal numbers wherever the channels are correlated)."""
    cfg = _abi.default_config(window_size=size[0], window_type=size[1], configuration=_abi.CH_PHASE, bin_polation=bp, **view)
    S, F, hop = 4, 4, size[0] // 4
    g = TransformPairBatch(cfg, S)
    assert g.info.fused_available == 1
    audio = frames_case(cfg, S, F, hop)
    o = OracleSpectrum(cfg, S)
    want = o.process_frames(audio, hop, F)
    got = g.process_frames(audio, hop, F)
    # the last display points scan the bins next to Nyquist, where the reference halves bin N/2 - 1 of channel 1 only:
    # on a flat noise floor the arg-max there flips with the FFT's rounding and the halved bin changes the value by a
    # quarter, so those two points are checked against the staged GPU path (same FFT) instead of the oracle
    compare_frames(got[..., :-2, :1], want[..., :-2, :1], cfg, "fused phase magnitudes")
    staged = TransformPairBatch(cfg, S)
    for f in range(F):
        assert staged.prepare_transform(audio[:, f * hop: f * hop + cfg.window_size])
        staged.do_transform(); staged.map_to_linear_space(); staged.post_process_std_transform()
    for k in range(2):
        sf, ss, so = g.get_states(k), staged.get_states(k), o.states(k)
        assert_linear_close(sf[..., 0], so[..., 0], frame_axes=(1,), what=f"phase states {k} vs oracle")
        assert_linear_close(sf[..., 0], ss[..., 0], frame_axes=(1,), what=f"phase states {k} vs staged")
        scale = so[..., 0].max(axis=1, keepdims=True)
        strong = so[..., 0] > 1e-3 * scale
        strong[:, -1] = False   # never-broke edge: the reference leaves the last cancellation stale
        assert np.all(np.abs(sf[..., 1] - so[..., 1])[strong] < 2e-3 * np.broadcast_to(scale, so[..., 1].shape)[strong])
        assert np.all(np.abs(sf[..., 1] - ss[..., 1])[strong] < 2e-3 * np.broadcast_to(scale, so[..., 1].shape)[strong])
        np.testing.assert_array_equal(g.get_results(k), got[:, -1, k])
        assert np.max(np.abs(g.get_results(k)[:, -2:, 0] - staged.get_results(k)[:, -2:, 0])) < 1e-4


@pytest.mark.parametrize("kw", [dict(), dict(window_size=4096, configuration=_abi.CH_LEFT, bin_polation=_abi.BIN_LINEAR),
                                dict(window_size=2048, configuration=_abi.CH_COMPLEX), dict(window_size=16384, configuration=_abi.CH_MIDSIDE, bin_polation=_abi.BIN_NONE),
                                dict(configuration=_abi.CH_PHASE), dict(window_size=4096, configuration=_abi.CH_PHASE, bin_polation=_abi.BIN_LINEAR)])
def test_frame_parallel_mode_gives_the_same_bits(kw):
    """few streams x many frames: the frames of a stream are spread over CTAs (parts that start from a zero state) and the
    peak hold is carried across the seams afterwards; results, states and the handle's last results must equal frame-by-
    frame calls (one frame per stream per launch: no seams) bit for bit, also across two calls (state carried)"""
    import torch
    cfg = _abi.default_config(**kw)
    S, F, hop = 2, 37, cfg.window_size // 8
    audio = frames_case(cfg, S, 2 * F, hop)
    first, second = audio[:, :signals.frames_length(cfg.window_size, hop, F)], audio[:, F * hop:]
    runs = {}
    for mode in ("parallel", "serial"):
        g = TransformPairBatch(cfg, S)
        d1, d2 = torch.from_numpy(np.ascontiguousarray(first)).cuda(), torch.from_numpy(np.ascontiguousarray(second)).cuda()
        if mode == "parallel":
            a = g.process_frames(d1, hop, F)   # device buffers: the call returns before the kernels have run, so the inputs stay alive
            b = g.process_frames(d2, hop, F)
        else:
            W = cfg.window_size   # (the slices stay alive until the handle's stream has been drained)
            s1 = [d1[:, f * hop:f * hop + W].contiguous() for f in range(F)]
            s2 = [d2[:, f * hop:f * hop + W].contiguous() for f in range(F)]
            torch.cuda.synchronize()
            a = [g.process_frames(x, hop, 1) for x in s1]
            b = [g.process_frames(x, hop, 1) for x in s2]
            g.synchronize()
            a, b = torch.cat(a, dim=1), torch.cat(b, dim=1)
        g.synchronize()
        a, b = a.cpu().numpy(), b.cpu().numpy()
        runs[mode] = (a, b, g.get_states(0), g.get_states(1), g.get_results(0), g.get_results(1))
    for x, y in zip(runs["parallel"], runs["serial"]):
        np.testing.assert_array_equal(x, y)
    want = OracleSpectrum(cfg, S).process_frames(first, hop, F)
    used = 2 if cfg.configuration in (_abi.CH_SEPARATE, _abi.CH_MIDSIDE) else 1
    last = -2 if cfg.configuration == _abi.CH_PHASE else None   # (Phase next to Nyquist: see test_fused_phase_configuration)
    compare_frames(runs["parallel"][0][..., :last, :used], want[..., :last, :used], cfg, "frame-parallel vs oracle")


@pytest.mark.parametrize("kw,S", [(dict(), 301), (dict(window_size=1024, bin_polation=_abi.BIN_LINEAR), 811),
                                  (dict(window_size=4096, configuration=_abi.CH_PHASE), 471), (dict(window_size=2048, configuration=_abi.CH_LEFT), 700)])
def test_balanced_ranges_chain_the_seams(kw, S):
    """more streams than resident CTAs: the launch gives every CTA the same number of frames, so most ranges end
    mid-stream; the head of such a stream runs first on one CTA, its state is handed to the next CTA, which runs the tail.
    Must equal frame-by-frame calls (whole streams per CTA, nothing to hand over) bit for bit, and the oracle."""
    import torch
    cfg = _abi.default_config(**kw)
    F, hop, W = 5, cfg.window_size // 4, cfg.window_size
    audio = frames_case(cfg, S, F, hop)
    d = torch.from_numpy(np.ascontiguousarray(audio)).cuda()
    g = TransformPairBatch(cfg, S)
    whole = g.process_frames(d, hop, F)
    g.synchronize()
    st = (g.get_states(0), g.get_states(1), g.get_results(0), g.get_results(1))
    g2 = TransformPairBatch(cfg, S)
    slices = [d[:, f * hop:f * hop + W].contiguous() for f in range(F)]
    torch.cuda.synchronize()
    steps = [g2.process_frames(x, hop, 1) for x in slices]
    g2.synchronize()
    steps = torch.cat(steps, dim=1)
    np.testing.assert_array_equal(whole.cpu().numpy(), steps.cpu().numpy())
    for x, y in zip(st, (g2.get_states(0), g2.get_states(1), g2.get_results(0), g2.get_results(1))):
        np.testing.assert_array_equal(x, y)
    sample = np.r_[0:3, S // 2:S // 2 + 3, S - 3:S]
    want = OracleSpectrum(cfg, len(sample)).process_frames(audio[sample], hop, F)
    used = 2 if cfg.configuration in (_abi.CH_SEPARATE, _abi.CH_MIDSIDE) else 1
    last = -2 if cfg.configuration == _abi.CH_PHASE else None
    compare_frames(whole.cpu().numpy()[sample][..., :last, :used], want[..., :last, :used], cfg, "chained seams vs oracle")


@pytest.mark.parametrize("kw,S,F", [(dict(), 150, 14), (dict(axis_points=777), 130, 11), (dict(window_size=4096, configuration=_abi.CH_PHASE), 200, 9),
                                    (dict(window_size=2048, configuration=_abi.CH_LEFT, bin_polation=_abi.BIN_LINEAR), 97, 23)])
def test_a_stream_in_two_or_three_parts(kw, S, F):
    """fewer streams than resident CTAs but not by much (the per-GPU shape of BASELINE config 3 on 4 or 8 GPUs): every
    stream is cut into two or three parts, k_seam_apply walks the parts before its own itself and whole-stream parts store
    their state directly.  Two calls in a row (the second starts from the first's states) against frame-by-frame calls,
    bit for bit."""
    import torch
    cfg = _abi.default_config(**kw)
    hop, W = cfg.window_size // 4, cfg.window_size
    audio = frames_case(cfg, S, 2 * F, hop)
    d = torch.from_numpy(np.ascontiguousarray(audio)).cuda()
    first, second = d[:, :signals.frames_length(W, hop, F)].contiguous(), d[:, F * hop:].contiguous()
    g = TransformPairBatch(cfg, S)
    a, b = g.process_frames(first, hop, F), g.process_frames(second, hop, F)
    g.synchronize()   # (the handle runs on its own stream: nothing may read the results before this)
    whole = torch.cat([a, b], dim=1)
    g2 = TransformPairBatch(cfg, S)
    slices = [d[:, f * hop:f * hop + W].contiguous() for f in range(2 * F)]
    torch.cuda.synchronize()
    steps = [g2.process_frames(x, hop, 1) for x in slices]
    g2.synchronize()
    np.testing.assert_array_equal(whole.cpu().numpy(), torch.cat(steps, dim=1).cpu().numpy())
    for k in (0, 1):
        np.testing.assert_array_equal(g.get_states(k), g2.get_states(k))
        np.testing.assert_array_equal(g.get_results(k), g2.get_results(k))


VIEWS = [dict(view_scaling=_abi.VIEW_LINEAR), dict(view_left=0.2, view_right=0.4), dict(view_scaling=_abi.VIEW_LINEAR, view_left=0.0, view_right=0.03),
         dict(view_left=0.6, view_right=1.0, low_dbs=-60.0, high_dbs=6.0), dict(view_scaling=_abi.VIEW_LINEAR, view_left=0.9, view_right=1.0, slope_a=0.5, slope_b=0.02)]


@pytest.mark.parametrize("view", VIEWS)
@pytest.mark.parametrize("bp", INTERPS)
@pytest.mark.parametrize("window", [8192, 4096, 2048, 1024])
def test_fused_views(view, bp, window):
    # zoomed / linear axes move the break between interpolated and max-of-bins points anywhere from "none" to "all",
    # which is what decides which shared-memory entries the fused kernels keep as magnitudes and which squared
    cfg = _abi.default_config(window_size=window, bin_polation=bp, **view)
    S, F, hop = 3, 3, window // 2
    g = TransformPairBatch(cfg, S)
    audio = frames_case(cfg, S, F, hop)
    got = g.process_frames(audio, hop, F)     # fused where the tables allow it, staged otherwise: same contract
    want = OracleSpectrum(cfg, S).process_frames(audio, hop, F)
    # per-stream peak of the whole spectrum on this view's display scale (slope at most b * (fs/2)^a), from a full-axis run
    full = _abi.default_config(window_size=window, bin_polation=_abi.BIN_NONE)
    st = OracleSpectrum(full, S); st.process_frames(audio, hop, F)
    slope_max = float(np.max(OracleSpectrum(cfg, 1).slope_map()))
    peak_floor = st.states(0)[..., :2].reshape(S, -1).max(axis=1).astype(np.float64) * slope_max / 0.95 ** F
    compare_frames(got, want, cfg, f"view {view} bp {bp} window {window} fused {g.info.fused_available}", peak_floor=peak_floor)


@pytest.mark.parametrize("window", [16384, 8192, 4096, 2048, 1024])
@pytest.mark.parametrize("points", [300, 777, 2048, 3840])
def test_fused_other_axis_sizes_and_hops_beyond_the_window(window, points):
    # display-point counts that are not a multiple of the CTA size, and frames that do not overlap at all
    cfg = _abi.default_config(window_size=window, axis_points=points)
    S, F, hop = 3, 3, window + 37
    g = TransformPairBatch(cfg, S)
    # fused as long as one CTA's shared memory (exchange buffer + 28 bytes per display point) fits an SM
    assert g.info.fused_available == int(window * 8 + 2048 + points * 28 <= 227 * 1024)
    audio = frames_case(cfg, S, F, hop)
    got = g.process_frames(audio, hop, F)
    want = OracleSpectrum(cfg, S).process_frames(audio, hop, F)
    compare_frames(got, want, cfg, f"window {window} points {points}")


@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
@pytest.mark.parametrize("hop", [2048, 1023, 1])
def test_fused_layouts_and_unaligned_hops(layout, hop):
    cfg = _abi.default_config()
    S, F = 3, 4
    audio = frames_case(cfg, S, F, hop, layout=layout)
    got = TransformPairBatch(cfg, S).process_frames(audio, hop, F, layout=layout)
    want = OracleSpectrum(cfg, S).process_frames(audio, hop, F, layout=layout)
    compare_frames(got, want, cfg, f"layout {layout} hop {hop}")


def test_fused_equals_staged_and_call_splitting_is_exact():
    """process_frames(F) == F x the four staged calls (same device arithmetic up to FFT
    algorithm), and splitting one call into two is bit-exact (state carried across calls)."""
    cfg = _abi.default_config()
    S, F, hop = 4, 5, 2048
    audio = frames_case(cfg, S, F, hop)
    a = TransformPairBatch(cfg, S)
    whole = a.process_frames(audio, hop, F)
    b = TransformPairBatch(cfg, S)
    part1 = b.process_frames(audio[:, :cfg.window_size + 2 * hop], hop, 3)
    part2 = b.process_frames(audio[:, 3 * hop:], hop, 2)
    np.testing.assert_array_equal(np.concatenate([part1, part2], axis=1), whole)
    c = TransformPairBatch(cfg, S)
    staged = np.empty_like(whole)
    for f in range(F):
        assert c.prepare_transform(audio[:, f * hop: f * hop + cfg.window_size])
        c.do_transform(); c.map_to_linear_space(); c.post_process_std_transform()
        staged[:, f, 0] = c.get_results(0); staged[:, f, 1] = c.get_results(1)
    compare_frames(whole, staged, cfg, "fused vs staged")


def test_device_resident_buffers_and_stream():
    import torch
    cfg = _abi.default_config()
    S, F, hop = 8, 3, 2048
    audio = frames_case(cfg, S, F, hop)
    host = TransformPairBatch(cfg, S).process_frames(audio, hop, F)
    g = TransformPairBatch(cfg, S)
    g.set_cuda_stream(torch.cuda.current_stream().cuda_stream)
    d_audio = torch.from_numpy(audio).cuda()
    d_out = g.process_frames(d_audio, hop, F)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(d_out.cpu().numpy(), host)


def test_staged_fallback_for_unfused_configs_in_process_frames():
    for conf, size in ((_abi.CH_PHASE, 512), (_abi.CH_COMPLEX, 512)):   # no fused kernel below 1024 points
        cfg = _abi.default_config(window_size=size, configuration=conf, axis_points=256)
        S, F, hop = 2, 3, size // 4
        g = TransformPairBatch(cfg, S)
        assert g.info.fused_available == 0
        audio = frames_case(cfg, S, F, hop)
        got = g.process_frames(audio, hop, F)
        want = OracleSpectrum(cfg, S).process_frames(audio, hop, F)
        compare_frames(got[..., :1], want[..., :1], cfg, f"staged frames conf {conf}")


def test_preliminary_audio_overload():
    # TransformDSP.inl:234-484 incl. its iterator quirk
    for conf in (_abi.CH_LEFT, _abi.CH_RIGHT, _abi.CH_MERGE, _abi.CH_SEPARATE, _abi.CH_MIDSIDE):
        cfg = _abi.default_config(window_size=2048, configuration=conf)
        S = 2
        hist = signals.batch(S, 2048)
        for n_prelim in (1, 200, 2048, 3000):
            prelim = signals.batch(S, n_prelim, first_stream=7)
            o = OracleSpectrum(cfg, S); g = TransformPairBatch(cfg, S)
            assert o.prepare_preliminary(hist, prelim) == _abi.OK
            assert g.prepare_transform(hist, preliminary=prelim)
            o.transform(); g.do_transform()
            complex_close(g.get_raw_fft(), o.raw_fft(), f"prelim conf {conf} n {n_prelim}", rel=2e-6, floor=2e-6)


def test_error_behaviour():
    cfg = _abi.default_config(window_size=4096)
    g = TransformPairBatch(cfg, 2)
    assert g.prepare_transform(signals.batch(2, 4095)) is False      # skip frame, TransformDSP.inl:45-46
    with pytest.raises(_lib.SgzError) as e:
        g.do_transform()                                               # call order, TransformPair.h:123-124
    assert e.value.code == _abi.ESTATE
    with pytest.raises(_lib.SgzError) as e:
        g.process_frames(signals.batch(2, 5000), 1024, 3)              # not enough samples
    assert e.value.code == _abi.EINVAL
    out = g.process_frames(signals.batch(2, 4096), 1024, 0)
    assert out.shape == (2, 0, 2, 1024, 2)
    g.clear_audio_state()
    assert np.all(g.get_states(0) == 0) and np.all(g.get_results(1) == 0)


def test_known_answers_on_gpu():
    # bin-centred sines read their amplitude; silence reads clipDB; scaling the input by 2 shifts
    # the display value by ln 2 / ln(high/low)
    cfg = _abi.default_config(bin_polation=_abi.BIN_NONE)
    W = 8192
    n = np.arange(W)
    x = np.zeros((3, W, 2), np.float32)
    x[0, :, 0] = 0.5 * np.sin(2 * np.pi * 1024 * n / W); x[0, :, 1] = 0.25 * np.cos(2 * np.pi * 1638 * n / W)
    x[2] = 2 * x[0]
    g = TransformPairBatch(cfg, 3)
    r = g.process_frames(x, W, 1)[:, 0, 0]
    assert abs(r[0, :, 0].max() - (1 - 6.0206 / 120)) < 1e-4
    assert abs(r[0, :, 1].max() - (1 - 12.0412 / 120)) < 1e-4
    assert np.all(r[1] == np.float32(-384.0))
    strong = r[0] > 0.5
    assert np.allclose((r[2] - r[0])[strong], np.log(2.0) / np.log(1e6), atol=2e-5)


def test_full_size_c3_properties():
    """BASELINE config 3 at full width (1024 stereo streams, N = 8192): batch independence (a
    stream's result does not depend on its batch), idempotence of peak hold on a repeated frame,
    and oracle parity on a sample of streams."""
    import torch
    cfg = _abi.default_config()
    S, F, hop = 1024, 4, 2048
    audio = frames_case(cfg, S, F, hop)
    g = TransformPairBatch(cfg, S)
    d_audio = torch.from_numpy(audio).cuda()
    d_got = g.process_frames(d_audio, hop, F)
    g.synchronize()  # device-buffer calls are asynchronous on the handle's stream
    got = d_got.cpu().numpy()
    assert np.isfinite(got).all()
    pick = [0, 1, 2, 511, 700, 1023]
    sub = np.ascontiguousarray(audio[pick])
    again = TransformPairBatch(cfg, len(pick)).process_frames(sub, hop, F)
    np.testing.assert_array_equal(again, got[pick])
    want = OracleSpectrum(cfg, len(pick)).process_frames(sub, hop, F, threads=4)
    compare_frames(again, want, cfg, "C3 sample")
    # idempotence: re-running the last frame only decays then re-peaks -> identical display values
    h = TransformPairBatch(cfg, len(pick))
    last = np.ascontiguousarray(sub[:, -cfg.window_size:])
    first = h.process_frames(last, hop, 1)
    second = h.process_frames(last, hop, 1)
    np.testing.assert_array_equal(first, second)


def test_bench_workload_against_the_oracle():
    """the workload bench.py times - 1024 streams x 64 frames, hop 2048, the same signals (bench.make_audio) - through
    the same device-resident call, compared with the oracle on 32 streams spread over the batch, every frame; the mean
    bench.py prints as `checksum` is recomputed from the oracle-checked values where the streams overlap"""
    import argparse
    import torch
    import bench
    args = argparse.Namespace(streams=1024, frames=64, hop=2048)
    audio, n = bench.make_audio(args, 0)
    cfg = bench.bench_cfg()
    S, F, hop = args.streams, args.frames, args.hop
    g = TransformPairBatch(cfg, S)
    d_audio = torch.from_numpy(audio).cuda()
    d_out = g.process_frames(d_audio, hop, F)
    g.synchronize()
    pick = np.unique(np.r_[np.arange(0, S, 37), 95, 96, 295, 296, S - 1])[:32]
    assert len(pick) == 32
    got = d_out[torch.from_numpy(pick).cuda()].cpu().numpy()
    want = OracleSpectrum(cfg, len(pick)).process_frames(np.ascontiguousarray(audio[pick]), hop, F, threads=8)
    assert np.isfinite(got).all()
    compare_frames(got, want, cfg, "bench workload")
    # the states the kernel leaves behind are the last frame's, whichever CTA ranges the streams fell into
    for k in range(2):
        st = g.get_states(k)[pick]
        lin = np.where(want[:, -1, k] > -300, 1e-6 * np.exp(want[:, -1, k].astype(np.float64) * np.log(1e6)), 0.0)
        assert_linear_close(st, lin, frame_axes=(1, 2), what=f"bench states {k}")


# ---------------------------------------------------------------------------------------------
# resonator bank (BASELINE config 2: RDFT mode, 1 stereo stream, 1024 log-spaced bins, 48 kHz)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("conf", [_abi.CH_LEFT, _abi.CH_MERGE, _abi.CH_SIDE, _abi.CH_PHASE, _abi.CH_SEPARATE, _abi.CH_MIDSIDE])
@pytest.mark.parametrize("win", [_abi.WIN_RECTANGULAR, _abi.WIN_HANN, _abi.WIN_BLACKMAN_HARRIS])
@pytest.mark.parametrize("S", [1, 3])
def test_resonator_matches_oracle(conf, win, S):
    cfg = _abi.default_config(algo=_abi.ALGO_RSNT, window_size=4096, window_type=win, configuration=conf)
    o = OracleSpectrum(cfg, S)
    g = TransformPairBatch(cfg, S)
    assert g.info.resonator_vectors == o.info.resonator_vectors
    pos = 0
    total = signals.batch(S, 60000)
    for block in (100, 1, 5000, 333, 48000):  # short blocks run sequentially, long ones are time-sliced
        chunk = np.ascontiguousarray(total[:, pos:pos + block]); pos += block
        o.resonate(chunk); g.resonating_dispatch(chunk)
        complex_close(g.get_resonator_state().reshape(S, -1, 2), o.resonator_state().reshape(S, -1, 2), f"state after {pos}", rel=1e-4, floor=2e-5)
    o.map(); g.map_to_linear_space()
    P, ch = cfg.axis_points, o.channels
    to = o.transform_result().reshape(S, ch, P, 2); tg = g.get_transform_result().reshape(S, ch, P, 2)
    if conf == _abi.CH_PHASE:
        assert_linear_close(tg[:, 0, :, 0], to[:, 0, :, 0], frame_axes=(1,), floor=2e-5, what="phase mid")
        strong = to[:, 0, :, 0] > 1e-2 * to[:, 0, :, 0].max(axis=1, keepdims=True)
        assert np.all(np.abs(tg[:, 0, :, 1][strong] - to[:, 0, :, 1][strong]) < 2e-3)
    else:
        complex_close(tg, to, "snapshot", rel=1e-4, floor=2e-5)
    o.postprocess(); g.post_process_std_transform()
    for k in range(2):
        so, sg = o.states(k), g.get_states(k)
        used = 1 if ch == 1 or conf == _abi.CH_PHASE else 2
        assert_linear_close(sg[..., :used], so[..., :used], frame_axes=(1, 2), floor=2e-5, what=f"peak-hold states {k}")


def test_resonator_frames_match_oracle_c2():
    # config 2 through the frame entry: `hop` samples of resonation per frame, then snapshot + post
    cfg = _abi.default_config(algo=_abi.ALGO_RSNT, window_size=8192, window_type=_abi.WIN_HANN, display_mode=_abi.DISPLAY_COLOURSPECTRUM,
                              sample_buffer_size=480)
    S, F, hop = 1, 20, 480
    audio = signals.batch(S, F * hop, first_stream=3)
    want = OracleSpectrum(cfg, S).process_frames(audio, hop, F)
    got = TransformPairBatch(cfg, S).process_frames(audio, hop, F)
    dy = 1.0 / np.log(1e6); low = 1e-6
    lin = np.where(want > -300, low * np.exp(want.astype(np.float64) / dy), 0.0)
    for f in range(F):
        m = want[0, f] > -300
        assert_log_close(got[0, f][m], want[0, f][m], lin=lin[0, f][m], peak=lin[0, f].max(), floor=2e-5, what=f"rsnt frame {f}")


def test_resonator_clear_state():
    cfg = _abi.default_config(algo=_abi.ALGO_RSNT, window_size=2048, window_type=_abi.WIN_HANN)
    g = TransformPairBatch(cfg, 2)
    g.resonating_dispatch(signals.batch(2, 1000))
    assert np.abs(g.get_resonator_state()).max() > 0
    g.clear_audio_state()
    assert np.all(g.get_resonator_state() == 0)


def test_host_buffer_pipeline_equals_single_launch():
    # >= 8 streams with host buffers take the chunked upload / compute / download pipeline
    import torch
    cfg = _abi.default_config()
    S, F, hop = 19, 3, 2048
    audio = frames_case(cfg, S, F, hop)
    a = TransformPairBatch(cfg, S)
    host = a.process_frames(audio, hop, F)                      # pipelined
    b = TransformPairBatch(cfg, S)
    dev = b.process_frames(torch.from_numpy(audio).cuda(), hop, F)
    b.synchronize()
    np.testing.assert_array_equal(host, dev.cpu().numpy())
    for k in range(2):
        np.testing.assert_array_equal(a.get_states(k), b.get_states(k))
        np.testing.assert_array_equal(a.get_results(k), b.get_results(k))
    again = a.process_frames(audio, hop, F)                     # state carried into a second pipelined call
    again_dev = b.process_frames(torch.from_numpy(audio).cuda(), hop, F)
    b.synchronize()
    np.testing.assert_array_equal(again, again_dev.cpu().numpy())


def test_spectrogram_colour_columns_bit_exact():
    # §8(f) row 1: blendAndDispatchSpectrums (SpectrumDSP.cpp:110-206) over process_frames results
    from oracle.binding import colour_columns, colour_ratios
    cfg = _abi.default_config(window_size=4096, window_type=_abi.WIN_HANN, display_mode=_abi.DISPLAY_COLOURSPECTRUM)
    S, F, hop, pairs = 6, 5, 1024, 3
    audio = frames_case(cfg, S, F, hop)
    g = TransformPairBatch(cfg, S)
    res = g.process_frames(audio, hop, F)
    ratios = colour_ratios([0.1, 0.2, 0.3, 0.25, 0.15])
    assert ratios[0] == 0 and abs(ratios.sum() - 1) < 1e-6
    rng = np.random.default_rng(5)
    colours = rng.random((pairs, 6, 3)).astype(np.float32)
    res[0, 0, 0, :7, 0] = [-1.0, 0.0, 0.9989, 0.999, 1.0, 1.5, 0.5]   # skipped, floor, near-top, top colour, ...
    got = g.blend_and_dispatch_spectrums(res, pairs, ratios, colours)
    want = colour_columns(res, pairs, ratios, colours)
    assert got.shape == (S // pairs, F, cfg.axis_points, 4)
    np.testing.assert_array_equal(got, want)
    assert np.all(got[..., 3] == 255) and got[..., :3].max() > 0


@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
@pytest.mark.parametrize("algo", [_abi.ALGO_FFT, _abi.ALGO_RSNT])
@pytest.mark.parametrize("conf", [_abi.CH_LEFT, _abi.CH_SEPARATE, _abi.CH_MIDSIDE])
def test_audio_entry_point_frame_scheduler(algo, conf, layout):
    # audioEntryPoint (TransformDSP.inl:1165-1211): blocks cut at sampleBufferSize boundaries, frames into sfbuf
    cfg = _abi.default_config(algo=algo, window_size=2048, window_type=_abi.WIN_HANN, configuration=conf,
                              display_mode=_abi.DISPLAY_COLOURSPECTRUM, sample_buffer_size=300, axis_points=256)
    S = 3
    o = OracleSpectrum(cfg, S)
    g = TransformPairBatch(cfg, S)
    hist = signals.batch(S, 2048, first_stream=1)
    total = signals.batch(S, 5000, first_stream=4)
    planar = layout == _abi.LAYOUT_PLANAR   # (planar: the right channel of a block starts at the block's own length)
    if planar:
        hist = np.ascontiguousarray(hist.transpose(0, 2, 1))
    pos = 0
    emitted = 0
    for block in (100, 250, 1024, 299, 301, 1500):   # boundaries fall inside, at the end of, and across blocks
        chunk = np.ascontiguousarray(total[:, pos:pos + block]); pos += block
        if planar:
            chunk = np.ascontiguousarray(chunk.transpose(0, 2, 1))
        want = o.audio_entry_point(hist, chunk, layout)
        got = g.audio_entry_point(hist if algo == _abi.ALGO_FFT else None, chunk, layout)
        assert got.shape == want.shape, (block, got.shape, want.shape)
        emitted += got.shape[1]
        if got.shape[1]:
            used = 1 if conf == _abi.CH_LEFT else 2
            lin = np.where(want > -300, 1e-6 * np.exp(want.astype(np.float64) * np.log(1e6)), 0.0)
            for s in range(S):
                for f in range(got.shape[1]):
                    m = want[s, f, :, :used] > -300
                    assert_log_close(got[s, f, :, :used][m], want[s, f, :, :used][m], lin=lin[s, f, :, :used][m], peak=lin[s, f].max(),
                                     floor=2e-5 if algo == _abi.ALGO_RSNT else FLOOR, what=f"block {block} frame {f}")
    assert emitted == sum((100, 250, 1024, 299, 301, 1500)) // 300
    # LineGraph mode: FFT does nothing here, RSNT resonates the whole block
    cfg2 = _abi.copy_config(cfg, display_mode=_abi.DISPLAY_LINEGRAPH)
    g2 = TransformPairBatch(cfg2, S)
    tail = total[:, :700].copy()
    assert g2.audio_entry_point(hist if algo == _abi.ALGO_FFT else None, np.ascontiguousarray(tail.transpose(0, 2, 1)) if planar else tail, layout).shape[1] == 0
    if algo == _abi.ALGO_RSNT:
        assert np.abs(g2.get_resonator_state()).max() > 0


@pytest.mark.parametrize("kw,S,host", [(dict(), 5, False), (dict(), 24, True), (dict(window_size=4096, configuration=_abi.CH_PHASE), 3, False),
                                        (dict(window_size=2048, configuration=_abi.CH_LEFT, bin_polation=_abi.BIN_NONE), 310, False),
                                        (dict(window_size=700, window_type=_abi.WIN_BLACKMAN, axis_points=300), 4, True),
                                        (dict(algo=_abi.ALGO_RSNT, window_size=2048, axis_points=128, window_type=_abi.WIN_HANN), 2, False)])
def test_output_selector_is_a_view_of_the_full_results(kw, S, host):
    """sgz_spectrum_process_frames_select: LineMain-of-every-frame and last-frame-only outputs are bit for bit the
    corresponding parts of the full output, states and lineGraphs results are the same afterwards - on the fused kernel
    (few streams: parts and seam fix-up; many streams: chained ranges; host buffers: the chunked pipeline) and on the
    staged / resonator paths"""
    import torch
    cfg = _abi.default_config(**kw)
    F, hop = 9, max(cfg.window_size // 4, 64)
    n = signals.frames_length(cfg.window_size, hop, F) if cfg.algo == _abi.ALGO_FFT else F * hop
    audio = signals.batch(S, n, first_stream=2)
    data = audio if host else torch.from_numpy(np.ascontiguousarray(audio)).cuda()
    runs = {}
    for mode in (_abi.OUT_FULL, _abi.OUT_LINE_MAIN, _abi.OUT_LAST_FRAME):
        g = TransformPairBatch(cfg, S)
        r = g.process_frames(data, hop, F, output=mode)
        g.synchronize()
        r = r if host else r.cpu().numpy()
        runs[mode] = (r, g.get_states(0), g.get_states(1), g.get_results(0), g.get_results(1))
    full = runs[_abi.OUT_FULL][0]
    np.testing.assert_array_equal(runs[_abi.OUT_LINE_MAIN][0], full[:, :, 0])
    np.testing.assert_array_equal(runs[_abi.OUT_LAST_FRAME][0], full[:, -1])
    np.testing.assert_array_equal(runs[_abi.OUT_FULL][3], full[:, -1, 0])     # lineGraphs[k].results = the last frame
    np.testing.assert_array_equal(runs[_abi.OUT_FULL][4], full[:, -1, 1])
    for mode in (_abi.OUT_LINE_MAIN, _abi.OUT_LAST_FRAME):
        for x, y in zip(runs[mode][1:], runs[_abi.OUT_FULL][1:]):
            np.testing.assert_array_equal(x, y)
