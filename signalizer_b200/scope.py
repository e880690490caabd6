"""Host-side mirror of the secondary views' hot arithmetic (Vectorscope, Oscilloscope) over the C ABI.

Reference call sites: `VectorScope::drawPolarPlot` / `drawRectPlot`
(Source/Vectorscope/VectorscopeRendering.cpp:444-746), `VectorScope::Processor::audioProcessing`
(Source/Vectorscope/Vectorscope.cpp:268-377), `Oscilloscope::drawWavePlot` Lanczos branch
(Source/Oscilloscope/OscilloscopeRendering.cpp:790-891), `ZeroCrossingProcessor` / `PeakHoldProcessor`
(Source/Oscilloscope/StreamPreprocessing.h:269-349) and `calculateFundamentalPeriod` /
`calculateTriggeringOffset` (Source/Oscilloscope/OscilloscopeDSP.inl:61-308)."""
import ctypes as C

import numpy as np

from . import _abi
from ._lib import check, load
from .spectrum import _buf


def _geometry(audio, layout):
    shape = tuple(audio.shape)
    assert len(shape) == 3
    n = shape[1] if layout == _abi.LAYOUT_INTERLEAVED else shape[2]
    return shape[0], n, shape[1] * shape[2]


class Context:
    def __init__(self, device: int = 0):
        self.lib = load()
        self._h = C.c_void_p()
        check(self.lib.sgz_context_create(int(device), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.sgz_context_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_cuda_stream(self, stream_ptr: int):
        check(self.lib.sgz_context_set_cuda_stream(self._h, C.c_void_p(stream_ptr)))

    def synchronize(self):
        check(self.lib.sgz_context_synchronize(self._h))

    # ---- Vectorscope ----------------------------------------------------------------------
    def vscope_vertices(self, cfg, audio, layout=_abi.LAYOUT_INTERLEAVED):
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        out = np.empty((S, n, 3), dtype=np.float32)
        check(self.lib.sgz_vscope_vertices(self._h, C.byref(cfg), ptr, mem, layout, S, n, stride, out.ctypes.data, _abi.MEM_HOST))
        return out

    def vscope_histogram(self, cfg, audio, layout=_abi.LAYOUT_INTERLEAVED, out=None):
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        if out is None:
            out = np.empty((S, cfg.height, cfg.width), dtype=np.uint32)
            check(self.lib.sgz_vscope_histogram(self._h, C.byref(cfg), ptr, mem, layout, S, n, stride, out.ctypes.data, _abi.MEM_HOST))
        else:  # device tensor (torch int32 view of the uint32 counts)
            check(self.lib.sgz_vscope_histogram(self._h, C.byref(cfg), ptr, mem, layout, S, n, stride, out.data_ptr(), _abi.MEM_DEVICE))
        return out

    def vscope_meters(self, audio, filters, envelope_coeff, stereo_coeff, second_speed, layout=_abi.LAYOUT_INTERLEAVED):
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        assert len(filters) == S
        check(self.lib.sgz_vscope_meters(self._h, ptr, mem, layout, S, n, stride, envelope_coeff, stereo_coeff, second_speed, filters))
        return filters

    def vscope_peak_envelope(self, audio, filters, gains, envelope_coeff, delta_time, layout=_abi.LAYOUT_INTERLEAVED):
        """VectorScope::runPeakFilter (VectorscopeRendering.cpp:826-889): folds the window's largest |L|, |R| into
        filters[s].envelope and returns the envelopeGain per stream (`gains` in, kept where the new value is not normal)."""
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        assert len(filters) == S
        g = np.ascontiguousarray(gains, dtype=np.float64).copy()
        assert g.shape == (S,)
        check(self.lib.sgz_vscope_peak_envelope(self._h, ptr, mem, layout, S, n, stride, envelope_coeff, delta_time, filters,
                                                g.ctypes.data_as(C.POINTER(C.c_double))))
        return filters, g

    # ---- Oscilloscope ---------------------------------------------------------------------
    def scope_resample_count(self, cfg) -> int:
        return int(self.lib.sgz_scope_resample_count(C.byref(cfg)))

    def scope_resample(self, cfg, sample_pos, audio, layout=_abi.LAYOUT_INTERLEAVED, out=None):
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        pos = np.ascontiguousarray(sample_pos, dtype=np.float64)
        assert pos.shape == (S,)
        count = self.scope_resample_count(cfg)
        if out is None:
            out = np.empty((S, count, 2), dtype=np.float32)
            optr, omem = out.ctypes.data, _abi.MEM_HOST
        else:
            optr, omem = out.data_ptr(), _abi.MEM_DEVICE
        check(self.lib.sgz_scope_resample(self._h, C.byref(cfg), pos.ctypes.data_as(C.POINTER(C.c_double)), ptr, mem, layout, S, n, stride, optr, omem))
        return out

    def scope_trigger(self, mode, channel, audio, states, threshold, hysteresis=0.0, steady_clock=0, max_peaks=4096, layout=_abi.LAYOUT_INTERLEAVED):
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        assert len(states) == S
        peaks = np.zeros((S, max_peaks), dtype=np.uint64)
        counts = np.zeros(S, dtype=np.uint32)
        check(self.lib.sgz_scope_trigger(self._h, mode, channel, ptr, mem, layout, S, n, stride, threshold, hysteresis, steady_clock, states,
                                         peaks.ctypes.data_as(C.POINTER(C.c_uint64)), counts.ctypes.data_as(C.POINTER(C.c_uint32)), max_peaks))
        return [peaks[s, :min(int(counts[s]), max_peaks)].copy() for s in range(S)], counts

    def scope_trigger_device(self, mode, channel, d_audio, d_states, d_peaks, d_counts, threshold, hysteresis=0.0, steady_clock=0, layout=_abi.LAYOUT_INTERLEAVED):
        """sgz_scope_trigger with every array a CUDA tensor (d_states: uint8 [S, 24] holding sgz_scope_trigger_state records,
        d_peaks: int64 [S, max_peaks], d_counts: int32 [S]); asynchronous on the context's stream."""
        keep, ptr, mem = _buf(d_audio)
        assert mem == _abi.MEM_DEVICE
        S, n, stride = _geometry(keep, layout)
        check(self.lib.sgz_scope_trigger_device(self._h, mode, channel, ptr, layout, S, n, stride, threshold, hysteresis, steady_clock,
                                                d_states.data_ptr(), d_peaks.data_ptr(), d_counts.data_ptr(), d_peaks.shape[1]))

    def scope_period(self, params, audio, states, layout=_abi.LAYOUT_INTERLEAVED):
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        assert len(states) == S
        results = (_abi.ScopePeriodResult * S)()
        check(self.lib.sgz_scope_period(self._h, C.byref(params), ptr, mem, layout, S, n, stride, states, results))
        return results

    def scope_colour(self, cfg, audio, states, layout=_abi.LAYOUT_INTERLEAVED, out=None):
        """Per-sample colouring + RMS envelope of StreamState::audioProcessing (OscilloscopeDSP.inl:426-710).
        states: (ScopeColourState * (2 * streams))(), in/out.  Returns (rgba [streams, 4, n, 4] uint8 with planes
        left, right, mid, side; envelope gains [streams])."""
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        assert len(states) == 2 * S
        gain = np.zeros(S, dtype=np.float64)
        if out is None:
            rgba = np.empty((S, 4, n, 4), dtype=np.uint8)
            optr, omem = rgba.ctypes.data, _abi.MEM_HOST
        else:   # a CUDA uint8 tensor [S, 4, n, 4]: the colour planes stay on the device
            rgba = out
            assert tuple(out.shape) == (S, 4, n, 4) and out.is_contiguous()
            optr, omem = out.data_ptr(), _abi.MEM_DEVICE
        check(self.lib.sgz_scope_colour(self._h, C.byref(cfg), ptr, mem, layout, S, n, stride, states, optr, omem,
                                        gain.ctypes.data_as(C.POINTER(C.c_double))))
        return rgba, gain

    def scope_peak_envelope(self, channel_mode, audio, envelopes, coeff, layout=_abi.LAYOUT_INTERLEAVED):
        """Oscilloscope::runPeakFilter (OscilloscopeDSP.inl:712-886): envelopes [streams, 2] float32 in/out; returns autoGain [streams]."""
        keep, ptr, mem = _buf(audio)
        S, n, stride = _geometry(keep, layout)
        assert envelopes.dtype == np.float32 and envelopes.shape == (S, 2) and envelopes.flags.c_contiguous
        gain = np.zeros(S, dtype=np.float64)
        check(self.lib.sgz_scope_peak_envelope(self._h, channel_mode, ptr, mem, layout, S, n, stride, coeff,
                                               envelopes.ctypes.data_as(C.POINTER(C.c_float)), gain.ctypes.data_as(C.POINTER(C.c_double))))
        return gain


def scope_trigger_windows(lib, states, peaks, peak_counts, steady_clock, n_samples, max_events=256):
    """TriggeringProcessor::update + processMutating (StreamPreprocessing.h:54-206) for every stream: consumes the queued
    trigger positions (`peaks` [S, max_peaks] uint64 with `peak_counts` valid entries per row, both updated in place) and
    returns the per-stream lists of (kind, origin, size, buffered) events.  Host-only bookkeeping: `lib` is the loaded
    library (signalizer_b200._lib.load() or the oracle's, which exports the same function under its own prefix)."""
    S = len(states)
    assert peaks.dtype == np.uint64 and peaks.shape[0] == S and peak_counts.dtype == np.uint32
    events = (_abi.ScopeWindowEvent * (S * max_events))()
    counts = np.zeros(S, dtype=np.uint32)
    fn = getattr(lib, "sgz_scope_trigger_windows", None) or lib.sgzo_scope_trigger_windows
    rc = fn(S, states, peaks.ctypes.data_as(C.POINTER(C.c_uint64)), peak_counts.ctypes.data_as(C.POINTER(C.c_uint32)), peaks.shape[1],
            steady_clock, n_samples, events, counts.ctypes.data_as(C.POINTER(C.c_uint32)), max_events)
    if rc != 0:
        raise ValueError(f"trigger_windows rc={rc}")
    return [[(e.kind, e.origin, e.size, e.buffered) for e in events[s * max_events:s * max_events + counts[s]]] for s in range(S)]
