"""Secondary paths on the GPU (through the C ABI) against the CPU oracle: Vectorscope vertices,
point-density histogram (counts bit-exact) and meters; Oscilloscope Lanczos resampler, stream
triggers (positions bit-exact) and the spectral period search."""
import numpy as np
import pytest

from oracle.binding import OracleScope
from signalizer_b200 import _abi, signals
from signalizer_b200.scope import Context

pytestmark = pytest.mark.gpu


def special_audio(S, n):
    """synthetic streams plus the edge cases of the polar transform: exact zeros, one silent
    channel, L == R, L == -R, full scale"""
    a = signals.batch(S, n)
    a[0, :64] = 0.0
    a[0, 64:128, 0] = 0.0
    a[0, 128:192, 1] = 0.0
    a[0, 192:256, 1] = a[0, 192:256, 0]
    a[0, 256:320, 1] = -a[0, 256:320, 0]
    a[0, 320:330] = 1.0
    a[0, 330:340] = -1.0
    return a


@pytest.mark.parametrize("polar,scale_polar,gain", [(True, False, 1.0), (True, True, 1.7), (False, False, 0.8), (True, False, 40.0)])
@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
def test_vectorscope_vertices_and_histogram_bit_exact(polar, scale_polar, gain, layout):
    S, n = 5, 30011
    a = special_audio(S, n)
    if layout == _abi.LAYOUT_PLANAR:
        a = np.ascontiguousarray(a.transpose(0, 2, 1))
    cfg = _abi.vscope_config(polar=polar, scale_polar=scale_polar, gain=gain, width=512, height=512)
    o, g = OracleScope(), Context()
    np.testing.assert_array_equal(g.vscope_vertices(cfg, a, layout), o.vscope_vertices(cfg, a, layout))
    hg, ho = g.vscope_histogram(cfg, a, layout), o.vscope_histogram(cfg, a, layout)
    np.testing.assert_array_equal(hg, ho)
    assert hg.sum() > 0
    if gain <= 1.0 and not scale_polar:
        assert np.all(hg.sum(axis=(1, 2)) + (np.abs(a).reshape(S, -1).max(axis=1) >= 1.0) * n >= n)  # nothing clipped unless full scale


def test_vectorscope_c5_full_size_histogram():
    # BASELINE config 5: 256 stereo streams, 512 x 512; device-resident buffers
    import torch
    S, n = 256, 48000
    base = signals.batch(32, n)
    a = np.ascontiguousarray(np.tile(base, (8, 1, 1)))
    cfg = _abi.vscope_config(gain=1.2)
    g = Context()
    d_a = torch.from_numpy(a).cuda()
    d_h = torch.empty((S, 512, 512), dtype=torch.int32, device="cuda")
    g.vscope_histogram(cfg, d_a, out=d_h)
    g.synchronize()
    h = d_h.cpu().numpy().view(np.uint32)
    want = OracleScope().vscope_histogram(cfg, a[:32])
    np.testing.assert_array_equal(h[:32], want)
    np.testing.assert_array_equal(h[32:64], want)   # tiled streams: identical counts
    assert int(h.sum()) <= S * n


def test_vectorscope_meters_bit_exact():
    S, n = 7, 12345
    a = special_audio(S, n)
    o, g = OracleScope(), Context()
    fo, fg = (_abi.VScopeFilters * S)(), (_abi.VScopeFilters * S)()
    for rep in range(2):  # state carried across calls
        o.vscope_meters(a, fo, 0.999, 0.9995, 5.0)
        g.vscope_meters(a, fg, 0.999, 0.9995, 5.0)
        np.testing.assert_array_equal(np.frombuffer(fg, dtype=np.float32), np.frombuffer(fo, dtype=np.float32))
    assert np.frombuffer(fg, dtype=np.float32).max() > 0


@pytest.mark.parametrize("channel", [_abi.OSC_LEFT, _abi.OSC_RIGHT, _abi.OSC_MID, _abi.OSC_SIDE])
@pytest.mark.parametrize("spp,inc", [(1.0 / 16, 1.0 / 2048), (0.3, 1.0 / 700), (1.0, 1.0 / 128)])
def test_scope_lanczos_resampler(channel, spp, inc):
    S, n = 6, 9000
    a = signals.batch(S, n)
    cfg = _abi.resample_config(channel=channel, samples_per_pixel=spp, left=0.0, right=1.0, inc=inc)
    pos = np.array([100.25, 77.0, 0.5, 3000.9, 512.125, 8980.0])  # the last one runs off the start of the buffer
    o, g = OracleScope(), Context()
    assert g.scope_resample_count(cfg) == o.scope_resample_count(cfg)
    want, got = o.scope_resample(cfg, pos, a), g.scope_resample(cfg, pos, a)
    np.testing.assert_array_equal(got[..., 0], want[..., 0])            # x positions: same double accumulation -> float
    assert np.max(np.abs(got[..., 1] - want[..., 1])) < 2e-6           # values: double sums rounded to float


@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
@pytest.mark.parametrize("ratio", [1, 2, 4, 64, 128, 3])
def test_scope_resampler_polyphase_path_edges(layout, ratio):
    # power-of-two oversampling ratios <= 64 take the polyphase kernel, the rest the direct one; a short
    # buffer makes most vertices touch the buffer edges, and the positions include exact sample alignment
    S, n = 5, 48
    a = signals.batch(S, n)
    buf = a if layout == _abi.LAYOUT_INTERLEAVED else np.ascontiguousarray(a.transpose(0, 2, 1))
    cfg = _abi.resample_config(channel=_abi.OSC_MID, samples_per_pixel=1.0 / ratio, left=-0.5, right=1.5, inc=1.0 / (ratio * 40))
    pos = np.array([30.0, 12.5, 47.0625, 0.0, 5.333])
    o, g = OracleScope(), Context()
    want, got = o.scope_resample(cfg, pos, buf, layout=layout), g.scope_resample(cfg, pos, buf, layout=layout)
    np.testing.assert_array_equal(got[..., 0], want[..., 0])
    assert np.max(np.abs(got[..., 1] - want[..., 1])) < 2e-6


def test_scope_c4_16x_resample_full_size():
    # BASELINE config 4: 256 streams, 16x windowed-sinc resampling; a band-limited sine is reproduced
    S, n = 256, 4096
    t = np.arange(n)
    a = np.zeros((S, n, 2), np.float32)
    freqs = 200.0 + 37.0 * np.arange(S)
    a[:, :, 0] = (0.5 * np.sin(2 * np.pi * freqs[:, None] * t[None, :] / 48000.0)).astype(np.float32)
    cfg = _abi.resample_config(channel=_abi.OSC_LEFT, samples_per_pixel=1.0 / 16, left=0.0, right=1.0, inc=1.0 / (16 * 2048))
    pos = np.full(S, 2100.0)
    got = Context().scope_resample(cfg, pos, a)
    assert got.shape == (S, 16 * 2048 + 1, 2)
    # output m sits at sample n - 2100 + m / 16 (m / 16 past kernel centre): compare with the analytic sine
    m = np.arange(got.shape[1])
    tt = (n - 2100) + m / 16.0
    want = 0.5 * np.sin(2 * np.pi * freqs[:, None] * tt[None, :] / 48000.0)
    assert np.max(np.abs(got[..., 1] - want)) < 2e-3   # Lanczos-10 passband ripple
    sub = OracleScope().scope_resample(cfg, pos[:4], a[:4])
    assert np.max(np.abs(got[:4, :, 1] - sub[..., 1])) < 2e-6


def test_zero_crossing_trigger_chunked_scan_equals_sequential():
    """the GPU cuts each stream into 16 chunks and chains their summaries: sparse, bursty and DC-shifted signals put
    the arming crossing, the firing sample and long event-free stretches on every side of the chunk borders"""
    rng = np.random.default_rng(11)
    o, g = OracleScope(), Context()
    for n in (1, 7, 32, 33, 511, 512, 513, 4096, 10007, 65536):
        S = 12
        a = np.zeros((S, n, 2), np.float32)
        for s in range(S):
            kind = s % 6
            if kind == 0:      # sparse spikes of either sign
                idx = rng.integers(0, n, size=max(1, n // 40)); a[s, idx, 0] = rng.uniform(-1, 1, size=len(idx))
            elif kind == 1:    # always above the threshold, never crossing: only the carried-in armed state can fire
                a[s, :, 0] = 0.5
            elif kind == 2:    # crossings but never above the threshold: stays armed through every chunk
                a[s, :, 0] = 0.01 * np.sign(np.sin(np.arange(n) * 0.37 + s))
            elif kind == 3:    # dense noise
                a[s, :, 0] = rng.uniform(-1, 1, size=n)
            elif kind == 4:    # one burst in the middle of silence
                m = n // 2; a[s, m: m + 5, 0] = [-0.3, 0.02, 0.04, 0.9, -0.2][: max(0, min(5, n - m))]
            else:              # slow sine: arming and firing far apart
                a[s, :, 0] = 0.8 * np.sin(np.arange(n) * 2 * np.pi / 3000.0 + s)
        so, sg = (_abi.ScopeTriggerState * S)(), (_abi.ScopeTriggerState * S)()
        for s in range(S):     # half of the streams come in armed
            for st in (so, sg):
                st[s].state = -0.25 if s % 2 else 0.3
                st[s].is_peak_holding = s % 2
                st[s].cross_origin = 1000 + s
        for cap in (4096, 3):
            po, co = o.scope_trigger(0, _abi.OSC_LEFT, a, so, threshold=0.05, steady_clock=5000, max_peaks=cap)
            pg, cg = g.scope_trigger(0, _abi.OSC_LEFT, a, sg, threshold=0.05, steady_clock=5000, max_peaks=cap)
            np.testing.assert_array_equal(cg, co)
            for s in range(S):
                np.testing.assert_array_equal(pg[s], po[s])
            assert bytes(sg) == bytes(so), n


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("channel", [_abi.OSC_LEFT, _abi.OSC_MID, _abi.OSC_SIDE])
def test_scope_triggers_bit_exact(mode, channel):
    S = 9
    total = signals.batch(S, 50000)
    o, g = OracleScope(), Context()
    so, sg = (_abi.ScopeTriggerState * S)(), (_abi.ScopeTriggerState * S)()
    clock = 1000
    pos = 0
    found = 0
    for block in (31, 1, 32, 4097, 20000, 25000):   # state carried across odd-sized blocks
        chunk = np.ascontiguousarray(total[:, pos:pos + block]); pos += block
        po, co = o.scope_trigger(mode, channel, chunk, so, threshold=0.05, hysteresis=0.3, steady_clock=clock)
        pg, cg = g.scope_trigger(mode, channel, chunk, sg, threshold=0.05, hysteresis=0.3, steady_clock=clock)
        np.testing.assert_array_equal(cg, co)
        for s in range(S):
            np.testing.assert_array_equal(pg[s], po[s])
            assert (sg[s].state, sg[s].is_peak_holding) == (so[s].state, so[s].is_peak_holding)
            if mode == 0:
                assert sg[s].cross_origin == so[s].cross_origin
        clock += block
        found += int(co.sum())
    assert found > 0


def test_scope_trigger_overflow_reports_full_count():
    a = signals.batch(2, 48000)
    g = Context()
    st = (_abi.ScopeTriggerState * 2)()
    peaks, counts = g.scope_trigger(0, _abi.OSC_LEFT, a, st, threshold=0.0, max_peaks=3)
    assert counts[0] > 3 and len(peaks[0]) == 3


def test_scope_spectral_period():
    S, n = 6, 12000
    t = np.arange(n)
    a = np.zeros((S, n, 2), np.float32)
    f0 = np.array([55.0, 110.0, 220.5, 441.0, 1000.0, 3000.0])
    for s in range(S):
        a[s, :, 0] = 0.5 * np.sin(2 * np.pi * f0[s] * t / 48000.0) + 0.2 * np.sin(2 * np.pi * 2 * f0[s] * t / 48000.0 + 0.3)
    prm = _abi.ScopePeriodParams(_abi.ABI_VERSION, _abi.OSC_LEFT, 2048.0, 48000.0, 0.01, 0.05, 0.0)
    o, g = OracleScope(), Context()
    so, sg = (_abi.ScopePeriodState * S)(), (_abi.ScopePeriodState * S)()
    for it in range(7):   # the 8-deep median needs five picks before it lets a new value through
        shift = 17 * it
        chunk = np.ascontiguousarray(a[:, :n - 200 + shift])
        ro, rg = o.scope_period(prm, chunk, so), g.scope_period(prm, chunk, sg)
        for s in range(S):
            assert rg[s].record.index == ro[s].record.index                     # bin index exact
            assert abs(rg[s].record.offset - ro[s].record.offset) < 1e-7
            assert abs(rg[s].fundamental - ro[s].fundamental) < 1e-6 * max(1.0, ro[s].fundamental)
            assert abs(rg[s].cycle_samples - ro[s].cycle_samples) < 1e-6 * ro[s].cycle_samples
            d = abs(rg[s].sample_offset - ro[s].sample_offset)
            assert min(d, abs(d - ro[s].cycle_samples)) < 1e-4                     # phase wraps at one cycle
            assert sg[s].median_pos == so[s].median_pos
    for s in range(S):
        assert abs(rg[s].fundamental - f0[s]) < 0.02 * f0[s]


@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
def test_vectorscope_peak_envelope_matches_oracle(layout):
    """VectorScope::runPeakFilter (VectorscopeRendering.cpp:826-889): envelope states and envelopeGain bit for bit, over
    several frames, incl. the whole-groups-of-8 rule (the loud sample in the remainder is not seen) and the isnormal guard
    (a silent stream keeps its previous gain)"""
    S, n = 7, 4099
    audio = signals.batch(S, n, first_stream=2)
    audio[3] = 0.0                      # silence: 1 / 0 is not normal, the gain stays
    audio[4, :, :] *= 0.01
    audio[4, n - 2, 0] = 0.9            # in the remainder (n & 7 = 3): ignored
    if layout == _abi.LAYOUT_PLANAR:
        audio = np.ascontiguousarray(audio.transpose(0, 2, 1))
    ctx, orc = Context(), OracleScope()
    fg, fo = (_abi.VScopeFilters * S)(), (_abi.VScopeFilters * S)()
    gg, go = np.full(S, 0.75), np.full(S, 0.75)
    for frame in range(3):
        scale = np.float32(1.0 / (1 + frame))
        fg, gg = ctx.vscope_peak_envelope(audio * scale, fg, gg, 0.97, 1.0 / 60.0, layout=layout)
        fo, go = orc.vscope_peak_envelope(audio * scale, fo, go, 0.97, 1.0 / 60.0, layout=layout)
        np.testing.assert_array_equal(gg, go)
        for s in range(S):
            assert tuple(fg[s].envelope) == tuple(fo[s].envelope)
    assert gg[3] == 0.75
    assert fg[4].envelope[0] < 0.01     # 0.9 was not seen


@pytest.mark.parametrize("mode", [0, 1])
def test_trigger_with_device_arrays_equals_host_arrays(mode):
    """sgz_scope_trigger_device: states, peak lists and counts stay on the GPU; same bits as the host-array call"""
    import ctypes as C
    import torch
    S, n, max_peaks = 9, 20000, 512
    audio = signals.batch(S, n, first_stream=1)
    ctx = Context()
    host_states = (_abi.ScopeTriggerState * S)()
    host_peaks, host_counts = ctx.scope_trigger(mode, _abi.OSC_MID, audio, host_states, 0.02, 0.1, steady_clock=777, max_peaks=max_peaks)
    d_audio = torch.from_numpy(audio).cuda()
    d_states = torch.zeros((S, C.sizeof(_abi.ScopeTriggerState)), dtype=torch.uint8, device="cuda")
    d_peaks = torch.zeros((S, max_peaks), dtype=torch.int64, device="cuda")
    d_counts = torch.zeros(S, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.scope_trigger_device(mode, _abi.OSC_MID, d_audio, d_states, d_peaks, d_counts, 0.02, 0.1, steady_clock=777)
    ctx.synchronize()
    counts = d_counts.cpu().numpy().astype(np.uint32)
    np.testing.assert_array_equal(counts, host_counts)
    peaks = d_peaks.cpu().numpy().astype(np.uint64)
    for s in range(S):
        np.testing.assert_array_equal(peaks[s, :min(int(counts[s]), max_peaks)], host_peaks[s])
    got_states = (_abi.ScopeTriggerState * S).from_buffer_copy(d_states.cpu().numpy().tobytes())
    for s in range(S):
        want = (host_states[s].state, host_states[s].cross_origin, host_states[s].is_peak_holding)
        assert (got_states[s].state, got_states[s].cross_origin, got_states[s].is_peak_holding) == want
