"""Host-side mirror of the reference's per-stream Spectrum interface, batched over streams.

`TransformPairBatch` exposes the members of `Signalizer::TransformPair<float>`
(Source/Spectrum/TransformPair.h:99-184) with the same names, call order
(prepare_transform -> do_transform -> map_to_linear_space -> post_process_std_transform,
TransformPair.h:123-124) and error behaviour (prepare_transform returns False to skip a frame,
TransformDSP.inl:45-46), over `n_streams` stereo pairs that share one `TransformConstant`
(SpectrumDSP.cpp:83-102).  All arithmetic runs in libsgz_cuda.so on the GPU; buffers may be numpy
arrays (host, copied inside the call) or torch CUDA tensors (used in place)."""
import ctypes as C

import numpy as np

from . import _abi
from ._lib import check, load


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _buf(x, dtype=np.float32, writable=False):
    """-> (keepalive, address, mem)"""
    if _is_torch(x):
        import torch
        assert x.dtype == torch.float32 and x.is_contiguous(), "float32 contiguous tensor required"
        return x, x.data_ptr(), (_abi.MEM_DEVICE if x.is_cuda else _abi.MEM_HOST)
    a = np.ascontiguousarray(x, dtype=dtype)
    if writable:
        assert a is x or np.shares_memory(a, x), "output array must be contiguous float32"
    return a, a.ctypes.data, _abi.MEM_HOST


class TransformConstant:
    """The read-only per-view constants (Source/Spectrum/TransformConstant.h:47-241), derived on
    the host by the library exactly when the reference's handleFlagUpdates would
    (Spectrum.cpp:351-616)."""

    def __init__(self, cfg: _abi.SpectrumConfig):
        self.cfg = _abi.copy_config(cfg)


class TransformPairBatch:
    def __init__(self, constant, n_streams: int = 1, device: int = 0):
        cfg = constant.cfg if isinstance(constant, TransformConstant) else constant
        self.lib = load()
        self.cfg = _abi.copy_config(cfg)
        self.n_streams = int(n_streams)
        self._h = C.c_void_p()
        check(self.lib.sgz_spectrum_create(C.byref(self.cfg), self.n_streams, int(device), C.byref(self._h)))
        self.info = _abi.SpectrumInfo()
        check(self.lib.sgz_spectrum_get_info(self._h, C.byref(self.info)))
        self.N = self.info.transform_size
        self.P = self.info.axis_points
        self.W = self.info.window_size
        self.channels = self.info.state_channels
        self.device = device  # -1 (SGZ_DEVICE_NONE) = tables-only handle, cannot compute

    # -- lifetime -------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.sgz_spectrum_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_cuda_stream(self, stream_ptr: int):
        check(self.lib.sgz_spectrum_set_cuda_stream(self._h, C.c_void_p(stream_ptr)))

    def synchronize(self):
        check(self.lib.sgz_spectrum_synchronize(self._h))

    # -- tables ---------------------------------------------------------------------------
    def _table(self, fn, n):
        out = np.empty(n, dtype=np.float32)
        check(fn(self._h, out.ctypes.data, n))
        return out

    def window_kernel(self):
        return self._table(self.lib.sgz_spectrum_get_window_kernel, self.N)

    def mapped_frequencies(self):
        return self._table(self.lib.sgz_spectrum_get_mapped_frequencies, self.P)

    def slope_map(self):
        return self._table(self.lib.sgz_spectrum_get_slope_map, self.P)

    def bin_map(self):
        arrs = [np.empty(self.P, dtype=np.int32) for _ in range(4)]
        check(self.lib.sgz_spectrum_get_bin_map(self._h, *[a.ctypes.data_as(C.POINTER(C.c_int32)) for a in arrs], self.P))
        return arrs

    # -- staged pipeline ------------------------------------------------------------------
    @staticmethod
    def _geometry(audio, layout):
        shape = tuple(audio.shape)
        assert len(shape) == 3, "audio must be [streams, samples, 2] (interleaved) or [streams, 2, samples] (planar)"
        n = shape[1] if layout == _abi.LAYOUT_INTERLEAVED else shape[2]
        return shape[0], n, shape[1] * shape[2]

    def prepare_transform(self, audio, layout=_abi.LAYOUT_INTERLEAVED, preliminary=None) -> bool:
        """bool prepareTransform(const Constant&, const AudioPair&[, preliminaryAudio, numSamples])
        (TransformDSP.inl:39-231, :234-484).  False = skip this frame."""
        keep, ptr, mem = _buf(audio)
        s, n, stride = self._geometry(keep, layout)
        assert s == self.n_streams
        if preliminary is None:
            rc = self.lib.sgz_spectrum_prepare(self._h, ptr, mem, layout, n, stride)
        else:
            keep2, ptr2, mem2 = _buf(preliminary)
            assert mem2 == mem, "history and preliminary audio must live in the same memory space"
            s2, n2, stride2 = self._geometry(keep2, layout)
            assert s2 == self.n_streams
            rc = self.lib.sgz_spectrum_prepare_preliminary(self._h, ptr, n, stride, ptr2, n2, stride2, mem, layout)
        check(rc, allow=(_abi.OK, _abi.SKIPPED))
        return rc == _abi.OK

    def do_transform(self):
        check(self.lib.sgz_spectrum_transform(self._h))

    def map_to_linear_space(self):
        check(self.lib.sgz_spectrum_map(self._h))

    def post_process_std_transform(self):
        check(self.lib.sgz_spectrum_postprocess(self._h))

    def get_raw_fft(self):
        out = np.empty((self.n_streams, self.N, 2), dtype=np.float32)
        check(self.lib.sgz_spectrum_get_raw_fft(self._h, out.ctypes.data, _abi.MEM_HOST))
        return out

    def get_transform_result(self):
        out = np.empty((self.n_streams, self.P * self.channels * 2), dtype=np.float32)
        check(self.lib.sgz_spectrum_get_transform_result(self._h, out.ctypes.data, _abi.MEM_HOST))
        return out

    def get_results(self, graph=_abi.LINE_MAIN):
        out = np.empty((self.n_streams, self.P, 2), dtype=np.float32)
        check(self.lib.sgz_spectrum_get_results(self._h, graph, out.ctypes.data, _abi.MEM_HOST))
        return out

    def get_states(self, graph=_abi.LINE_MAIN):
        out = np.empty((self.n_streams, self.P, 2), dtype=np.float32)
        check(self.lib.sgz_spectrum_get_states(self._h, graph, out.ctypes.data, _abi.MEM_HOST))
        return out

    def set_states(self, graph, states):
        keep, ptr, mem = _buf(states)
        assert tuple(keep.shape) == (self.n_streams, self.P, 2)
        check(self.lib.sgz_spectrum_set_states(self._h, graph, ptr, mem))

    def clear_line_graph_states(self):
        check(self.lib.sgz_spectrum_clear_line_graph_states(self._h))

    def clear_audio_state(self):
        check(self.lib.sgz_spectrum_clear_audio_state(self._h))

    # -- audioEntryPoint ---------------------------------------------------------------------
    def audio_entry_point(self, history, block, layout=_abi.LAYOUT_INTERLEAVED, max_frames=64):
        """template<ISA> void audioEntryPoint(const Constant&, const optional<AudioPair>&, array<AFloat*, 2>, size_t)
        (TransformDSP.inl:1165-1211) for every pair; returns the emitted sfbuf frames [streams, n, P, 2]."""
        bk, bptr, mem = _buf(block)
        s, nb, bstride = self._geometry(bk, layout)
        assert s == self.n_streams
        if history is not None:
            hk, hptr, hmem = _buf(history)
            assert hmem == mem
            _, nh, hstride = self._geometry(hk, layout)
        else:
            hptr, nh, hstride = None, 0, 0
        out = np.zeros((self.n_streams, max_frames, self.P, 2), dtype=np.float32)
        n = C.c_uint32(0)
        check(self.lib.sgz_spectrum_audio_entry_point(self._h, hptr, nh, hstride, bptr, nb, bstride, mem, layout, out.ctypes.data, max_frames, C.byref(n)))
        return out[:, :min(n.value, max_frames)]

    # -- spectrogram colour columns ----------------------------------------------------------
    def blend_and_dispatch_spectrums(self, results, pairs_per_view, ratios, colours):
        """AudioDispatcher::blendAndDispatchSpectrums (SpectrumDSP.cpp:110-206): results
        [streams, frames, 2, P, 2] -> RGBA8 columns [streams / pairs_per_view, frames, P, 4]."""
        keep, ptr, mem = _buf(results)
        n_frames = keep.shape[1]
        ratios = np.ascontiguousarray(ratios, dtype=np.float32); colours = np.ascontiguousarray(colours, dtype=np.float32)
        assert ratios.shape == (6,) and colours.shape == (pairs_per_view, 6, 3)
        out = np.empty((self.n_streams // pairs_per_view, n_frames, self.P, 4), dtype=np.uint8)
        check(self.lib.sgz_spectrum_colour_columns(self._h, ptr, mem, n_frames, pairs_per_view, ratios.ctypes.data, colours.ctypes.data,
                                                   out.ctypes.data, _abi.MEM_HOST))
        return out

    # -- resonator bank (algo = RSNT) ------------------------------------------------------
    def set_resonator_path(self, path: int) -> None:
        """0 = tensor cores for blocks of at least 8192 samples (default), 1 = the FP32 recurrence always"""
        check(self.lib.sgz_resonator_set_path(self._h, path))

    def resonating_dispatch(self, audio, layout=_abi.LAYOUT_INTERLEAVED):
        """template<ISA> void resonatingDispatch(const Constant&, array<AFloat*, 2> buffer, size_t n)
        (TransformDSP.inl:1215-1295) for every stream."""
        keep, ptr, mem = _buf(audio)
        s, n, stride = self._geometry(keep, layout)
        assert s == self.n_streams
        check(self.lib.sgz_resonator_process(self._h, ptr, mem, layout, n, stride))

    def get_resonator_state(self):
        V = self.info.resonator_vectors
        out = np.empty((self.n_streams, self.channels, V, self.P, 2), dtype=np.float32)
        check(self.lib.sgz_resonator_get_state(self._h, out.ctypes.data, _abi.MEM_HOST))
        return out

    # -- fused batched path ---------------------------------------------------------------
    def reserve(self, n_samples: int, stream_stride: int, n_frames: int) -> None:
        """size every lazily grown buffer once (staging, pipeline streams / events, resonator scratch)"""
        check(self.lib.sgz_spectrum_reserve(self._h, n_samples, stream_stride, n_frames))

    def kernel_name(self) -> str:
        return self.lib.sgz_spectrum_kernel_name(self._h).decode()

    def process_frames(self, audio, hop: int, n_frames: int, layout=_abi.LAYOUT_INTERLEAVED, out=None, output=_abi.OUT_FULL):
        """For every stream, `n_frames` frames at `hop`: the loop body of
        SpectrumRendering.cpp:622-634 / addAudioFrame (TransformDSP.inl:1139-1148).
        Returns results [streams, frames, 2 graphs, P, 2]; with output=OUT_LINE_MAIN only graph LineMain of every frame
        ([streams, frames, P, 2], what addAudioFrame keeps), with OUT_LAST_FRAME only the last frame ([streams, 2, P, 2])."""
        keep, ptr, mem = _buf(audio)
        s, n, stride = self._geometry(keep, layout)
        assert s == self.n_streams
        shape = {_abi.OUT_FULL: (self.n_streams, n_frames, 2, self.P, 2), _abi.OUT_LINE_MAIN: (self.n_streams, n_frames, self.P, 2),
                 _abi.OUT_LAST_FRAME: (self.n_streams, 2, self.P, 2)}[output]
        if out is None:
            if mem == _abi.MEM_DEVICE:
                import torch
                out = torch.empty(shape, dtype=torch.float32, device=keep.device)
            else:
                out = np.empty(shape, dtype=np.float32)
        okeep, optr, omem = _buf(out, writable=True)
        assert tuple(okeep.shape) == shape
        check(self.lib.sgz_spectrum_process_frames_select(self._h, ptr, mem, layout, n, stride, hop, n_frames, optr, omem, output))
        return out

    # -- device-resident audio history -------------------------------------
    def history_reserve(self, capacity: int) -> None:
        """Allocate (or clear) the per-stream history ring: the device-side stand-in for the
        cpl::AudioStream history the reference reads at TransformDSP.inl:49,65-88."""
        check(self.lib.sgz_spectrum_history_reserve(self._h, capacity))

    def history_append(self, block, layout=_abi.LAYOUT_INTERLEAVED) -> None:
        """Append one callback block ([streams, n, 2] or planar [streams, 2, n])."""
        keep, ptr, mem = _buf(block)
        s, n, stride = self._geometry(keep, layout)
        assert s == self.n_streams
        check(self.lib.sgz_spectrum_history_append(self._h, ptr, mem, layout, n, stride))

    def history_info(self):
        """(capacity, total samples appended, position of the last processed frame boundary)"""
        c, t, f = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(self.lib.sgz_spectrum_history_info(self._h, C.byref(c), C.byref(t), C.byref(f)))
        return c.value, t.value, f.value

    def history_read(self, n_last: int) -> np.ndarray:
        """The last `n_last` samples of every stream, oldest first, [streams, n_last, 2]."""
        out = np.empty((self.n_streams, n_last, 2), dtype=np.float32)
        check(self.lib.sgz_spectrum_history_read(self._h, n_last, out.ctypes.data, _abi.MEM_HOST))
        return out

    def history_view(self, n_last: int):
        """(device address, stream stride in floats) of the last `n_last` samples as one contiguous
        interleaved run per stream; valid until the next append."""
        p, stride = C.c_void_p(), C.c_uint64()
        check(self.lib.sgz_spectrum_history_view(self._h, n_last, C.byref(p), C.byref(stride)))
        return p.value, stride.value

    def prepare_transform_device_view(self, ptr: int, n: int, stride: int) -> bool:
        """prepareTransform on a raw device view (as returned by history_view)."""
        rc = self.lib.sgz_spectrum_prepare(self._h, ptr, _abi.MEM_DEVICE, _abi.LAYOUT_INTERLEAVED, n, stride)
        check(rc, allow=(_abi.OK, _abi.SKIPPED))
        return rc == _abi.OK

    def history_process(self, hop: int, max_frames: int, out=None, device=False):
        """Run every frame that fell due since the last call (at most `max_frames`); returns
        (results[:, :n], n) with results laid out like process_frames."""
        shape = (self.n_streams, max_frames, 2, self.P, 2)
        if out is None:
            if device:
                import torch
                out = torch.empty(shape, dtype=torch.float32, device=f"cuda:{self.device}")
            else:
                out = np.empty(shape, dtype=np.float32)
        okeep, optr, omem = _buf(out, writable=True)
        n = C.c_uint32(0)
        check(self.lib.sgz_spectrum_history_process(self._h, hop, max_frames, optr, omem, C.byref(n)))
        # the library packs [stream][n][graph][P][2] with n, not max_frames, as the frame stride
        flat = out.reshape(-1)
        got = flat[: self.n_streams * n.value * 2 * self.P * 2].reshape(self.n_streams, n.value, 2, self.P, 2)
        return got, n.value

    # -- frequency tracker --------------------------------------------------
    def track_peaks(self, graph: int, mouse_fraction, smoothing_ms: float = 0.0, fps: float = 60.0, states=None):
        """The peak search of Spectrum::drawFrequencyTracking (SpectrumRendering.cpp:301-469) for every
        stream: returns a ctypes array of TrackerResult.  `states` = (TrackerState * streams)() to
        run SmoothedPeakState (Spectrum.h:405-459) as well."""
        mf = np.ascontiguousarray(np.broadcast_to(np.asarray(mouse_fraction, dtype=np.float64), (self.n_streams,)))
        out = (_abi.TrackerResult * self.n_streams)()
        check(self.lib.sgz_spectrum_track_peaks(self._h, graph, mf.ctypes.data_as(C.POINTER(C.c_double)), smoothing_ms, fps, states, out))
        return out


def fft_forward_noalloc(d_in, d_out, n: int, batch: int, cuda_stream: int = 0) -> None:
    """Forward FFT of `batch` complex sequences of `n` points on DEVICE tensors, for callers that may
    not allocate GPU memory when they call (real-time callbacks, hosts that forbid cudaMalloc after
    start-up): wraps sgz_fft_forward_noalloc, which owns a fixed 8 MiB working buffer inside the
    library instead of allocating one.  `d_in` / `d_out` are distinct float32 CUDA tensors holding
    batch * n interleaved (re, im) pairs; n * batch <= 2**20.  Asynchronous on `cuda_stream`;
    calls that overlap on the GPU share (and corrupt) the one working buffer."""
    lib = load()
    for t in (d_in, d_out):
        assert _is_torch(t) and t.is_cuda and t.is_contiguous(), "device float32 tensors required"
        assert t.numel() >= 2 * n * batch
    check(lib.sgz_fft_forward_noalloc(C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), n, batch, C.c_void_p(cuda_stream)))
