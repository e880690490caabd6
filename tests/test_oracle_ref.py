"""Pins the restated oracle (oracle/sgz_oracle.hpp) with the REFERENCE's own source.

oracle/_ref/libsgz_oracle_ref.so (oracle/build_ref.sh) is Source/Spectrum/{TransformDSP.inl, TransformPair.h,
TransformConstant.h} compiled unmodified from the reference checkout against stand-in cpl headers whose primitives
(FFT, interpolation filters, window, resonator) forward to the oracle's "cpl decisions".  Every statement of
prepareTransform / doTransform / mapToLinearSpace / postProcessStdTransform / audioEntryPoint / resonatingDispatch /
remapFrequencies / generateSlopeMap / setStorage that runs there is the reference's, so bit-for-bit agreement here means
the restatement follows the in-tree arithmetic exactly; the cpl primitives remain decisions.

The library is built where the reference checkout is (this container) and travels to the GPU box prebuilt; where it
is missing altogether these tests skip."""
import numpy as np
import pytest

from oracle import binding
from oracle.binding import OracleSpectrum
from signalizer_b200 import _abi, signals

binding.build()
pytestmark = pytest.mark.skipif(not binding.ref_available(), reason="oracle/_ref not built (no reference checkout)")

CONFS = list(range(8))
INTERPS = [_abi.BIN_NONE, _abi.BIN_LINEAR, _abi.BIN_LANCZOS]
# window size, window, display points: transform sizes 4096, 8192, 1024, 32, 32768, 512, 16384, 65536
SIZES = [(4096, _abi.WIN_HANN, 1024), (8192, _abi.WIN_BLACKMAN_HARRIS, 1024), (700, _abi.WIN_BLACKMAN, 300), (20, _abi.WIN_HANN, 40),
         (20000, _abi.WIN_NUTTALL, 512), (400, _abi.WIN_HANN, 128), (10000, _abi.WIN_HANN, 512), (40000, _abi.WIN_BLACKMAN_HARRIS, 700)]
VIEWS = [dict(), dict(view_scaling=_abi.VIEW_LINEAR), dict(view_left=0.2, view_right=0.4), dict(view_scaling=_abi.VIEW_LINEAR, view_left=0.0, view_right=0.03),
         dict(view_left=0.6, view_right=1.0, low_dbs=-60.0, high_dbs=6.0, slope_a=0.5, slope_b=0.02)]


def both(cfg, S):
    return OracleSpectrum(cfg, S), OracleSpectrum(cfg, S, ref=True)


def staged(o, audio):
    """the render-thread sequence, every observable in between"""
    out = [o.window_kernel(), o.mapped_frequencies(), o.slope_map()]
    assert o.prepare(audio) == 0
    out.append(o.raw_fft())
    o.transform(); out.append(o.raw_fft())
    o.map(); out.append(o.transform_result())
    o.postprocess()
    out += [o.results(0), o.results(1), o.states(0), o.states(1)]
    return out


@pytest.mark.parametrize("size", SIZES)
@pytest.mark.parametrize("bp", INTERPS)
@pytest.mark.parametrize("conf", CONFS)
def test_staged_pipeline_bit_for_bit(conf, bp, size):
    W, win, P = size
    cfg = _abi.default_config(window_size=W, window_type=win, axis_points=P, configuration=conf, bin_polation=bp)
    S = 2
    audio = signals.batch(S, W + 37, first_stream=conf)
    a, b = both(cfg, S)
    for k in range(2):   # second pass: the peak-hold states carry over
        for x, y in zip(staged(a, audio), staged(b, audio)):
            np.testing.assert_array_equal(x, y)
        audio = audio * np.float32(0.5)


@pytest.mark.parametrize("view", VIEWS)
@pytest.mark.parametrize("bp", INTERPS)
@pytest.mark.parametrize("conf", [_abi.CH_LEFT, _abi.CH_PHASE, _abi.CH_SEPARATE, _abi.CH_COMPLEX])
def test_views_and_frames_bit_for_bit(conf, bp, view):
    cfg = _abi.default_config(window_size=2048, configuration=conf, bin_polation=bp, axis_points=384, **view)
    S, F, hop = 2, 4, 512
    audio = signals.batch(S, signals.frames_length(cfg.window_size, hop, F), first_stream=3)
    a, b = both(cfg, S)
    np.testing.assert_array_equal(a.process_frames(audio, hop, F), b.process_frames(audio, hop, F))
    for k in range(2):
        np.testing.assert_array_equal(a.states(k), b.states(k))


@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
@pytest.mark.parametrize("conf", CONFS)
def test_preliminary_audio_overload_bit_for_bit(conf, layout):
    """prepareTransform(constant, views, preliminaryAudio, numSamples), TransformDSP.inl:234-484, incl. ring splits"""
    cfg = _abi.default_config(window_size=1500, configuration=conf, axis_points=200)
    S = 2
    hist = signals.batch(S, 2000, first_stream=1)
    prelim = signals.batch(S, 333, first_stream=5)
    if layout == _abi.LAYOUT_PLANAR:
        hist, prelim = np.ascontiguousarray(hist.transpose(0, 2, 1)), np.ascontiguousarray(prelim.transpose(0, 2, 1))
    for split in (None, 0, 700, 1990):
        a, b = both(cfg, S)
        assert a.prepare_preliminary(hist, prelim, layout, split) == b.prepare_preliminary(hist, prelim, layout, split)
        np.testing.assert_array_equal(a.raw_fft(), b.raw_fft())


@pytest.mark.parametrize("algo", [_abi.ALGO_FFT, _abi.ALGO_RSNT])
@pytest.mark.parametrize("conf", [_abi.CH_LEFT, _abi.CH_MERGE, _abi.CH_SIDE, _abi.CH_PHASE, _abi.CH_SEPARATE, _abi.CH_MIDSIDE, _abi.CH_COMPLEX])
def test_audio_entry_point_bit_for_bit(conf, algo):
    """the frame scheduler of the colour-spectrogram mode, TransformDSP.inl:1165-1211: blocks that straddle blob boundaries"""
    cfg = _abi.default_config(window_size=1024, configuration=conf, algo=algo, axis_points=128, window_type=_abi.WIN_HANN,
                              display_mode=_abi.DISPLAY_COLOURSPECTRUM, sample_buffer_size=200)
    S = 2
    hist = signals.batch(S, 4096, first_stream=2)
    a, b = both(cfg, S)
    for n in (64, 200, 513, 199, 1, 450):
        block = signals.batch(S, n, first_stream=7 + n)
        fa, fb = a.audio_entry_point(hist, block), b.audio_entry_point(hist, block)
        assert fa.shape == fb.shape
        np.testing.assert_array_equal(fa, fb)
    if algo == _abi.ALGO_RSNT:
        np.testing.assert_array_equal(a.resonator_state(), b.resonator_state())


@pytest.mark.parametrize("conf", CONFS)
def test_resonator_path_bit_for_bit(conf):
    """resonatingDispatch (TransformDSP.inl:1215-1295) + the RSNT branch of mapToLinearSpace (:1103-1133)"""
    cfg = _abi.default_config(window_size=2048, configuration=conf, algo=_abi.ALGO_RSNT, axis_points=96, window_type=_abi.WIN_HANN)
    S = 2
    audio = signals.batch(S, 3000, first_stream=4)
    a, b = both(cfg, S)
    for o in (a, b):
        o.resonate(audio)
    np.testing.assert_array_equal(a.resonator_state(), b.resonator_state())
    for o in (a, b):
        o.map(); o.postprocess()
    np.testing.assert_array_equal(a.transform_result(), b.transform_result())
    for k in range(2):
        np.testing.assert_array_equal(a.results(k), b.results(k))


def test_skip_frame_and_clear_calls_agree():
    cfg = _abi.default_config(window_size=512, axis_points=64)
    a, b = both(cfg, 1)
    short = signals.batch(1, 100)
    assert a.prepare(short) == b.prepare(short) == _abi.SKIPPED   # TransformDSP.inl:45-46
    audio = signals.batch(1, 512)
    for o in (a, b):
        o.prepare(audio); o.transform(); o.map(); o.postprocess(); o.clear_line_graph_states()
    np.testing.assert_array_equal(a.states(0), b.states(0))
    for o in (a, b):
        o.clear_audio_state()
    np.testing.assert_array_equal(a.raw_fft(), b.raw_fft())
