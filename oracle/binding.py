"""ctypes binding of the CPU oracle (oracle/libsgz_oracle*.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never import this from signalizer_b200/.
Pinned by the reference's own transform source (oracle/_ref, tests/test_oracle_ref.py); cpl primitives are
decisions — see oracle/sgz_oracle.hpp."""
import ctypes as C
import os
import subprocess

import numpy as np

from signalizer_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


REF_LIB = os.path.join(_HERE, "_ref", "libsgz_oracle_ref.so")


def build(force: bool = False) -> None:
    subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    # oracle/_ref: the reference's own TransformDSP.inl behind the same entry points; needs the reference checkout
    # (this container), a no-op elsewhere (the GPU box uses the prebuilt file that travels with the snapshot)
    subprocess.check_call(["bash", os.path.join(_HERE, "build_ref.sh")])


def ref_available() -> bool:
    return os.path.exists(REF_LIB)


def _cpu_has_avx2() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            txt = f.read()
        return " avx2" in txt and " fma" in txt
    except OSError:
        return False


def load(fast: bool = False, ref: bool = False):
    """fast: the -O3 -mavx2 build of the restatement (timing only); ref: oracle/_ref, the reference's own
    TransformDSP.inl compiled against the cpl stand-ins (same entry points, Spectrum path only)."""
    quick = fast and _cpu_has_avx2()
    name = ("_ref/libsgz_oracle_ref_fast.so" if quick else "_ref/libsgz_oracle_ref.so") if ref else "libsgz_oracle_fast.so" if quick else "libsgz_oracle.so"
    if name in _libs:
        return _libs[name]
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        build()
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path}: oracle/_ref can only be built where the reference checkout is (oracle/build_ref.sh)")
    lib = C.CDLL(path)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    fp, dp, ip = C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int32)
    lib.sgzo_create.restype = vp
    lib.sgzo_create.argtypes = [C.POINTER(_abi.SpectrumConfig), u32, i32]
    lib.sgzo_destroy.argtypes = [vp]
    lib.sgzo_get_info.argtypes = [vp, C.POINTER(_abi.SpectrumInfo)]
    for n in ("sgzo_get_window_kernel", "sgzo_get_mapped_frequencies", "sgzo_get_slope_map",
              "sgzo_get_raw_fft", "sgzo_get_transform_result"):
        getattr(lib, n).argtypes = [vp, dp]
    lib.sgzo_get_bin_map.argtypes = [vp, ip, ip, ip, ip]
    lib.sgzo_prepare.argtypes = [vp, fp, i32, u64, u64, u64]
    lib.sgzo_prepare_preliminary.argtypes = [vp, fp, u64, u64, u64, fp, u64, u64, i32]
    for n in ("sgzo_transform", "sgzo_map", "sgzo_postprocess", "sgzo_clear_line_graph_states", "sgzo_clear_audio_state"):
        getattr(lib, n).argtypes = [vp]
    lib.sgzo_get_results.argtypes = [vp, i32, fp]
    lib.sgzo_get_states.argtypes = [vp, i32, fp]
    lib.sgzo_set_states.argtypes = [vp, i32, fp]
    lib.sgzo_process_frames.argtypes = [vp, fp, i32, u64, u64, u32, u32, fp, i32, u32, u32]
    lib.sgzo_resonate.argtypes = [vp, fp, i32, u64, u64]
    lib.sgzo_get_resonator_state.argtypes = [vp, dp]
    lib.sgzo_audio_entry_point.argtypes = [vp, fp, u64, u64, fp, u64, u64, i32, fp, u32, C.POINTER(u32)]
    lib.sgzo_hardware_concurrency.restype = C.c_uint
    _libs[name] = lib
    return lib


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


class OracleSpectrum:
    """`n_streams` TransformPair objects sharing one TransformConstant (float32 like the
    reference, Spectrum.h:71; `double=True` instantiates the same restatement in float64)."""

    def __init__(self, cfg: _abi.SpectrumConfig, n_streams: int = 1, double: bool = False, fast: bool = False, ref: bool = False):
        self.lib = load(fast, ref)
        self.cfg = _abi.copy_config(cfg)
        self.n_streams = n_streams
        self.h = self.lib.sgzo_create(C.byref(self.cfg), n_streams, int(double))
        if not self.h:
            raise ValueError("sgzo_create rejected the configuration")
        self.info = _abi.SpectrumInfo()
        self.lib.sgzo_get_info(self.h, C.byref(self.info))
        self.N = self.info.transform_size
        self.P = self.info.axis_points
        self.channels = self.info.state_channels

    def close(self):
        if self.h:
            self.lib.sgzo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dbl(self, fn, n):
        out = np.empty(n, dtype=np.float64)
        fn(self.h, out.ctypes.data_as(C.POINTER(C.c_double)))
        return out

    def window_kernel(self):
        return self._dbl(self.lib.sgzo_get_window_kernel, self.N)

    def mapped_frequencies(self):
        return self._dbl(self.lib.sgzo_get_mapped_frequencies, self.P)

    def slope_map(self):
        return self._dbl(self.lib.sgzo_get_slope_map, self.P)

    def bin_map(self):
        arrs = [np.empty(self.P, dtype=np.int32) for _ in range(4)]
        self.lib.sgzo_get_bin_map(self.h, *[a.ctypes.data_as(C.POINTER(C.c_int32)) for a in arrs])
        return arrs

    @staticmethod
    def _layout(audio):
        # [S, n, 2] interleaved or [S, 2, n] planar is ambiguous by shape alone; callers pass layout
        raise NotImplementedError

    def prepare(self, audio, layout=_abi.LAYOUT_INTERLEAVED, split=None):
        a, p = _f32(audio)
        n = a.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else a.shape[2]
        stride = a[0].size
        return self.lib.sgzo_prepare(self.h, p, layout, n, stride, n if split is None else split)

    def prepare_preliminary(self, history, prelim, layout=_abi.LAYOUT_INTERLEAVED, split=None):
        h, hp = _f32(history)
        q, qp = _f32(prelim)
        nh = h.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else h.shape[2]
        nq = q.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else q.shape[2]
        return self.lib.sgzo_prepare_preliminary(self.h, hp, nh, h[0].size, nh if split is None else split, qp, nq, q[0].size, layout)

    def transform(self):
        self.lib.sgzo_transform(self.h)

    def map(self):
        self.lib.sgzo_map(self.h)

    def postprocess(self):
        self.lib.sgzo_postprocess(self.h)

    def raw_fft(self):
        out = self._dbl(self.lib.sgzo_get_raw_fft, self.n_streams * self.N * 2)
        return out.reshape(self.n_streams, self.N, 2)

    def transform_result(self):
        n = self.P * self.channels * 2
        return self._dbl(self.lib.sgzo_get_transform_result, self.n_streams * n).reshape(self.n_streams, n)

    def results(self, graph=0):
        out = np.empty((self.n_streams, self.P, 2), dtype=np.float32)
        self.lib.sgzo_get_results(self.h, graph, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def states(self, graph=0):
        out = np.empty((self.n_streams, self.P, 2), dtype=np.float32)
        self.lib.sgzo_get_states(self.h, graph, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def set_states(self, graph, states):
        a, p = _f32(states)
        assert a.shape == (self.n_streams, self.P, 2)
        self.lib.sgzo_set_states(self.h, graph, p)

    def clear_line_graph_states(self):
        self.lib.sgzo_clear_line_graph_states(self.h)

    def clear_audio_state(self):
        self.lib.sgzo_clear_audio_state(self.h)

    def process_frames(self, audio, hop, n_frames, layout=_abi.LAYOUT_INTERLEAVED, threads=1,
                       want_results=True, first_stream=0, num_streams=None):
        a, p = _f32(audio)
        n = a.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else a.shape[2]
        ns = self.n_streams - first_stream if num_streams is None else num_streams
        out = np.zeros((self.n_streams, n_frames, 2, self.P, 2), dtype=np.float32) if want_results else None
        op = out.ctypes.data_as(C.POINTER(C.c_float)) if want_results else None
        rc = self.lib.sgzo_process_frames(self.h, p, layout, n, a[0].size, hop, n_frames, op, threads, first_stream, ns)
        if rc != 0:
            raise ValueError(f"sgzo_process_frames rc={rc}")
        return out

    def resonate(self, audio, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        n = a.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else a.shape[2]
        return self.lib.sgzo_resonate(self.h, p, layout, n, a[0].size)

    def resonator_state(self):
        V = self.info.resonator_vectors
        out = self._dbl(self.lib.sgzo_get_resonator_state, self.n_streams * self.channels * V * self.P * 2)
        return out.reshape(self.n_streams, self.channels, V, self.P, 2)

    def audio_entry_point(self, history, block, layout=_abi.LAYOUT_INTERLEAVED, max_frames=64):
        h, hp = _f32(history)
        b, bp = _f32(block)
        nh = h.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else h.shape[2]
        nb = b.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else b.shape[2]
        out = np.zeros((self.n_streams, max_frames, self.P, 2), dtype=np.float32)
        nf = C.c_uint32(0)
        self.lib.sgzo_audio_entry_point(self.h, hp, nh, h[0].size, bp, nb, b[0].size, layout,
                                        out.ctypes.data_as(C.POINTER(C.c_float)), max_frames, C.byref(nf))
        return out[:, :nf.value]


def hardware_concurrency() -> int:
    return int(load().sgzo_hardware_concurrency())


# ---- secondary paths (oracle/sgz_oracle_scope.hpp) -------------------------------------------------
def _scope_lib():
    lib = load()
    if getattr(lib, "_scope_ready", False):
        return lib
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    fp = C.POINTER(C.c_float)
    vc = C.POINTER(_abi.VScopeConfig)
    lib.sgzo_vscope_vertices.argtypes = [vc, fp, i32, u32, u64, u64, fp]
    lib.sgzo_vscope_histogram.argtypes = [vc, fp, i32, u32, u64, u64, C.POINTER(C.c_uint32)]
    lib.sgzo_vscope_meters.argtypes = [fp, i32, u32, u64, u64, C.c_float, C.c_float, C.c_float, C.POINTER(_abi.VScopeFilters)]
    rc = C.POINTER(_abi.ScopeResampleConfig)
    lib.sgzo_scope_resample_count.argtypes = [rc]
    lib.sgzo_scope_resample_count.restype = C.c_uint64
    lib.sgzo_scope_resample.argtypes = [rc, C.POINTER(C.c_double), fp, i32, u32, u64, u64, fp]
    lib.sgzo_scope_trigger.argtypes = [i32, i32, fp, i32, u32, u64, u64, C.c_double, C.c_double, u64, C.POINTER(_abi.ScopeTriggerState),
                                       C.POINTER(C.c_uint64), C.POINTER(u32), u32]
    lib.sgzo_scope_period.argtypes = [C.POINTER(_abi.ScopePeriodParams), fp, i32, u32, u64, u64, C.POINTER(_abi.ScopePeriodState),
                                      C.POINTER(_abi.ScopePeriodResult)]
    lib._scope_ready = True
    return lib


def _geom(a, layout):
    n = a.shape[1] if layout == _abi.LAYOUT_INTERLEAVED else a.shape[2]
    return a.shape[0], n, a[0].size


class OracleScope:
    """CPU restatement of the Vectorscope / Oscilloscope arithmetic (test infrastructure only)."""

    def __init__(self):
        self.lib = _scope_lib()

    def vscope_vertices(self, cfg, audio, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        out = np.empty((S, n, 3), dtype=np.float32)
        self.lib.sgzo_vscope_vertices(C.byref(cfg), p, layout, S, n, stride, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def vscope_histogram(self, cfg, audio, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        out = np.empty((S, cfg.height, cfg.width), dtype=np.uint32)
        self.lib.sgzo_vscope_histogram(C.byref(cfg), p, layout, S, n, stride, out.ctypes.data_as(C.POINTER(C.c_uint32)))
        return out

    def vscope_meters(self, audio, filters, envelope_coeff, stereo_coeff, second_speed, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        self.lib.sgzo_vscope_meters(p, layout, S, n, stride, envelope_coeff, stereo_coeff, second_speed, filters)
        return filters

    def vscope_peak_envelope(self, audio, filters, gains, envelope_coeff, delta_time, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        g = np.ascontiguousarray(gains, dtype=np.float64)
        self.lib.sgzo_vscope_peak_envelope.restype = None
        self.lib.sgzo_vscope_peak_envelope.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_uint64, C.c_float, C.c_double,
                                                       C.c_void_p, C.POINTER(C.c_double)]
        self.lib.sgzo_vscope_peak_envelope(p, layout, S, n, stride, envelope_coeff, delta_time, C.addressof(filters), g.ctypes.data_as(C.POINTER(C.c_double)))
        return filters, g

    def scope_resample_count(self, cfg):
        return int(self.lib.sgzo_scope_resample_count(C.byref(cfg)))

    def scope_resample(self, cfg, sample_pos, audio, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        pos = np.ascontiguousarray(sample_pos, dtype=np.float64)
        out = np.empty((S, self.scope_resample_count(cfg), 2), dtype=np.float32)
        self.lib.sgzo_scope_resample(C.byref(cfg), pos.ctypes.data_as(C.POINTER(C.c_double)), p, layout, S, n, stride, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def scope_colour(self, cfg, audio, states, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        rgba = np.empty((S, 4, n, 4), dtype=np.uint8)
        gain = np.zeros(S, dtype=np.float64)
        self.lib.sgzo_scope_colour.restype = None
        self.lib.sgzo_scope_colour.argtypes = [C.POINTER(_abi.ScopeColourConfig), C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_uint64,
                                               C.POINTER(_abi.ScopeColourState), C.c_void_p, C.POINTER(C.c_double)]
        self.lib.sgzo_scope_colour(C.byref(cfg), p, layout, S, n, stride, states, rgba.ctypes.data, gain.ctypes.data_as(C.POINTER(C.c_double)))
        return rgba, gain

    def scope_peak_envelope(self, channel_mode, audio, envelopes, coeff, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        gain = np.zeros(S, dtype=np.float64)
        self.lib.sgzo_scope_peak_envelope.restype = None
        self.lib.sgzo_scope_peak_envelope.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_uint64, C.c_double, C.c_void_p, C.POINTER(C.c_double)]
        self.lib.sgzo_scope_peak_envelope(channel_mode, p, layout, S, n, stride, coeff, envelopes.ctypes.data, gain.ctypes.data_as(C.POINTER(C.c_double)))
        return gain

    def scope_trigger(self, mode, channel, audio, states, threshold, hysteresis=0.0, steady_clock=0, max_peaks=4096, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        peaks = np.zeros((S, max_peaks), dtype=np.uint64)
        counts = np.zeros(S, dtype=np.uint32)
        self.lib.sgzo_scope_trigger(mode, channel, p, layout, S, n, stride, threshold, hysteresis, steady_clock, states,
                                    peaks.ctypes.data_as(C.POINTER(C.c_uint64)), counts.ctypes.data_as(C.POINTER(C.c_uint32)), max_peaks)
        return [peaks[s, :min(int(counts[s]), max_peaks)].copy() for s in range(S)], counts

    def scope_period(self, params, audio, states, layout=_abi.LAYOUT_INTERLEAVED):
        a, p = _f32(audio)
        S, n, stride = _geom(a, layout)
        results = (_abi.ScopePeriodResult * S)()
        self.lib.sgzo_scope_period(C.byref(params), p, layout, S, n, stride, states, results)
        return results


def colour_ratios(spec_ratios):
    lib = load()
    lib.sgzo_colour_ratios.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_float)]
    a = np.ascontiguousarray(spec_ratios, dtype=np.float64)
    out = np.empty(6, dtype=np.float32)
    lib.sgzo_colour_ratios(a.ctypes.data_as(C.POINTER(C.c_double)), out.ctypes.data_as(C.POINTER(C.c_float)))
    return out


def colour_columns(results, pairs, ratios, colours):
    lib = load()
    lib.sgzo_colour_columns.argtypes = [C.POINTER(C.c_float), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_float),
                                        C.POINTER(C.c_float), C.POINTER(C.c_uint8)]
    r, rp = _f32(results)
    S, F, _, P, _ = r.shape
    ra, rap = _f32(ratios)
    co, cop = _f32(colours)
    out = np.empty((S // pairs, F, P, 4), dtype=np.uint8)
    lib.sgzo_colour_columns(rp, S, F, P, pairs, rap, cop, out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out


def track_peaks(cfg, graph, transform_size, window_scale, mapped_freq, slope_map, mouse_fraction, results=None, raw_fft=None,
                smoothing_ms=0.0, fps=60.0, states=None):
    """oracle/sgz_oracle_tracker.hpp on caller-supplied arrays: results [S, P, 2] float32 or raw_fft
    [S, N + 1, 2] float32 (the searched array is handed in so that both sides search the same data)."""
    lib = load()
    src = results if results is not None else raw_fft
    S = src.shape[0]
    mf = np.ascontiguousarray(np.broadcast_to(np.asarray(mouse_fraction, dtype=np.float64), (S,)))
    f = np.ascontiguousarray(mapped_freq, dtype=np.float32); sl = np.ascontiguousarray(slope_map, dtype=np.float32)
    out = (_abi.TrackerResult * S)()
    fp = C.POINTER(C.c_float)
    lib.sgzo_track_peaks.restype = None
    lib.sgzo_track_peaks.argtypes = [C.POINTER(_abi.SpectrumConfig), C.c_int, C.c_uint32, C.c_uint32, C.c_double, fp, fp, fp, fp, C.POINTER(C.c_double),
                                     C.c_double, C.c_double, C.POINTER(_abi.TrackerState), C.POINTER(_abi.TrackerResult)]
    r = np.ascontiguousarray(results, dtype=np.float32) if results is not None else None
    w = np.ascontiguousarray(raw_fft, dtype=np.float32) if raw_fft is not None else None
    lib.sgzo_track_peaks(C.byref(cfg), graph, S, transform_size, window_scale, f.ctypes.data_as(fp), sl.ctypes.data_as(fp),
                         r.ctypes.data_as(fp) if r is not None else None, w.ctypes.data_as(fp) if w is not None else None,
                         mf.ctypes.data_as(C.POINTER(C.c_double)), smoothing_ms, fps, states, out)
    return out
