"""Oscilloscope colouring + envelopes through the C ABI (sgz_scope_colour, sgz_scope_peak_envelope)
against the oracle restatement of StreamState::audioProcessing / runPeakFilter
(OscilloscopeDSP.inl:426-886): every RGBA byte, every filter-state float and both envelopes bit-exact
(the kernel uses explicitly rounded float operations in the oracle's order)."""
import numpy as np
import pytest

from oracle.binding import OracleScope
from signalizer_b200 import _abi, _lib, signals
from signalizer_b200.scope import Context

pytestmark = pytest.mark.gpu
MODES = [_abi.OSC_LEFT, _abi.OSC_RIGHT, _abi.OSC_MID, _abi.OSC_SIDE, _abi.OSC_SEPARATE, _abi.OSC_MIDSIDE]


def states(S):
    return (_abi.ScopeColourState * (2 * S))()


def special(S, n):
    a = signals.batch(S, n)
    a[0, : n // 3] = 0.0                 # leading silence: 0 * inf colours
    a[1, :, 1] = a[1, :, 0]              # identical channels: side bands exactly zero
    if S > 2:
        a[2, :, 1] = 0.0
    return a


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
def test_colour_streams_and_states_bit_exact_across_blocks(mode, layout):
    S = 37                               # odd: one idle lane pair in the last warp
    total = special(S, 5000)
    cfg = _abi.colour_config(channel_mode=mode, envelope_mode=_abi.ENVELOPE_RMS, colour_smoothing_ms=2.5, envelope_window=0.02, colouring_blend=0.7,
                             colours=((0.9, 0.1, 0.0), (0.2, 1.0, 0.3), (0.0, 0.25, 0.95)), key_left=(12, 200, 34, 255), key_right=(250, 3, 99, 128))
    o, g = OracleScope(), Context()
    so, sg = states(S), states(S)
    pos = 0
    for n in (1, 3, 4, 999, 2048, 1945):  # odd block sizes: tails of the 4-pixel stores, state carried across calls
        blk = np.ascontiguousarray(total[:, pos:pos + n]); pos += n
        if layout == _abi.LAYOUT_PLANAR:
            blk = np.ascontiguousarray(blk.transpose(0, 2, 1))
        want, gw = o.scope_colour(cfg, blk, so, layout=layout)
        got, gg = g.scope_colour(cfg, blk, sg, layout=layout)
        np.testing.assert_array_equal(got, want)
        assert bytes(sg) == bytes(so)
        np.testing.assert_array_equal(gg, gw)
    assert len(np.unique(got.reshape(-1, 4), axis=0)) > 50   # the colours actually vary


def test_colour_many_streams_path_bit_exact():
    # from 2048 pairs on the library switches to the packed kernel (16 pairs per warp)
    S, n = 2051, 301
    a = np.tile(special(7, n), (293, 1, 1))[:S].copy()
    a *= np.linspace(0.1, 1.0, S, dtype=np.float32)[:, None, None]
    cfg = _abi.colour_config(channel_mode=_abi.OSC_MIDSIDE, envelope_mode=_abi.ENVELOPE_RMS, colouring_blend=0.9)
    o, g = OracleScope(), Context()
    so, sg = states(S), states(S)
    for blk in (a[:, :150].copy(), a[:, 150:].copy()):
        want, gw = o.scope_colour(cfg, blk, so)
        got, gg = g.scope_colour(cfg, blk, sg)
        np.testing.assert_array_equal(got, want)
        assert bytes(sg) == bytes(so)
        np.testing.assert_array_equal(gg, gw)


def test_colour_envelope_modes_and_errors():
    S, n = 4, 1024
    a = signals.batch(S, n)
    o, g = OracleScope(), Context()
    for env_mode in (_abi.ENVELOPE_NONE, _abi.ENVELOPE_PEAK):
        cfg = _abi.colour_config(channel_mode=_abi.OSC_SEPARATE, envelope_mode=env_mode)
        so, sg = states(S), states(S)
        want, gw = o.scope_colour(cfg, a, so)
        got, gg = g.scope_colour(cfg, a, sg)
        np.testing.assert_array_equal(got, want)
        assert bytes(sg) == bytes(so) and np.all(gg == 0) and all(sg[i].envelope == 0 for i in range(2 * S))
    bad = _abi.colour_config(channel_mode=9)
    with pytest.raises(_lib.SgzError):
        g.scope_colour(bad, a, states(S))


@pytest.mark.parametrize("mode", MODES)
def test_peak_envelope_bit_exact(mode):
    S = 9
    a = signals.batch(S, 4099)           # 3 samples beyond the last whole 8-vector
    a[3] *= 0.01
    o, g = OracleScope(), Context()
    eo = np.full((S, 2), 0.05, np.float32); eg = eo.copy()
    for coeff in (0.93, 0.5, 1.0):
        gw = o.scope_peak_envelope(mode, a, eo, coeff)
        gg = g.scope_peak_envelope(mode, a, eg, coeff)
        np.testing.assert_array_equal(eg, eo)
        np.testing.assert_array_equal(gg, gw)
    assert eg.max() > 0.05
