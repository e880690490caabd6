"""Loader for the CUDA shared library (signalizer_b200/libsgz_cuda.so).

The product path has no CPU fallback: if the library is missing this raises, and every compute
entry point returns SGZ_ECUDA (raised as SgzError) when no CUDA device is present."""
import ctypes as C
import os
import subprocess

from . import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# SGZ_CUDA_LIB: development hook, loads an A/B build of the same sources (csrc/Makefile VARIANT=...) instead
LIB_PATH = os.environ.get("SGZ_CUDA_LIB") or os.path.join(_HERE, "libsgz_cuda.so")
_lib = None

# every symbol include/sgz.h declares
EXPORTS = [
    "sgz_last_error", "sgz_spectrum_default_config", "sgz_spectrum_create", "sgz_spectrum_destroy",
    "sgz_spectrum_get_info", "sgz_spectrum_set_cuda_stream", "sgz_spectrum_synchronize",
    "sgz_spectrum_get_window_kernel", "sgz_spectrum_get_mapped_frequencies", "sgz_spectrum_get_slope_map",
    "sgz_spectrum_get_bin_map", "sgz_spectrum_prepare", "sgz_spectrum_prepare_preliminary",
    "sgz_spectrum_transform", "sgz_spectrum_map", "sgz_spectrum_postprocess", "sgz_spectrum_get_raw_fft",
    "sgz_spectrum_get_transform_result", "sgz_spectrum_get_results", "sgz_spectrum_get_states",
    "sgz_spectrum_set_states", "sgz_spectrum_clear_line_graph_states", "sgz_spectrum_clear_audio_state",
    "sgz_spectrum_process_frames", "sgz_spectrum_process_frames_select", "sgz_fft_forward_noalloc",
    "sgz_spectrum_history_reserve", "sgz_spectrum_history_append", "sgz_spectrum_history_info", "sgz_spectrum_history_view",
    "sgz_spectrum_history_read", "sgz_spectrum_history_process", "sgz_spectrum_track_peaks",
    "sgz_resonator_process", "sgz_resonator_get_state", "sgz_spectrum_colour_columns", "sgz_spectrum_audio_entry_point",
    "sgz_context_create", "sgz_context_destroy", "sgz_context_set_cuda_stream", "sgz_context_synchronize",
    "sgz_vscope_vertices", "sgz_vscope_histogram", "sgz_vscope_meters",
    "sgz_scope_resample_count", "sgz_scope_resample", "sgz_scope_trigger", "sgz_scope_period", "sgz_scope_colour", "sgz_scope_peak_envelope",
    "sgz_vscope_peak_envelope", "sgz_scope_trigger_windows", "sgz_spectrum_kernel_name", "sgz_scope_trigger_device", "sgz_resonator_set_path", "sgz_spectrum_reserve",
]


class SgzError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sgz error {code}: {msg}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile the library for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    csrc = os.path.join(_HERE, "csrc")
    cmd = ["make", "-C", csrc, "-s", "-j4"] + (["-B"] if force else [])
    subprocess.check_call(cmd)
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, u32, u64, i32, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_size_t
    fp, ip = C.c_void_p, C.POINTER(C.c_int32)  # float buffers are passed as raw addresses (host or device)
    lib.sgz_last_error.restype = C.c_char_p
    lib.sgz_spectrum_default_config.argtypes = [C.POINTER(_abi.SpectrumConfig)]
    lib.sgz_spectrum_default_config.restype = None
    lib.sgz_spectrum_create.argtypes = [C.POINTER(_abi.SpectrumConfig), u32, i32, C.POINTER(vp)]
    lib.sgz_spectrum_destroy.argtypes = [vp]
    lib.sgz_spectrum_destroy.restype = None
    lib.sgz_spectrum_get_info.argtypes = [vp, C.POINTER(_abi.SpectrumInfo)]
    lib.sgz_spectrum_set_cuda_stream.argtypes = [vp, vp]
    lib.sgz_spectrum_kernel_name.argtypes = [vp]
    lib.sgz_spectrum_kernel_name.restype = C.c_char_p
    lib.sgz_spectrum_synchronize.argtypes = [vp]
    for n in ("sgz_spectrum_get_window_kernel", "sgz_spectrum_get_mapped_frequencies", "sgz_spectrum_get_slope_map"):
        getattr(lib, n).argtypes = [vp, fp, sz]
    lib.sgz_spectrum_get_bin_map.argtypes = [vp, ip, ip, ip, ip, sz]
    lib.sgz_spectrum_prepare.argtypes = [vp, fp, i32, i32, u64, u64]
    lib.sgz_spectrum_prepare_preliminary.argtypes = [vp, fp, u64, u64, fp, u64, u64, i32, i32]
    for n in ("sgz_spectrum_transform", "sgz_spectrum_map", "sgz_spectrum_postprocess",
              "sgz_spectrum_clear_line_graph_states", "sgz_spectrum_clear_audio_state"):
        getattr(lib, n).argtypes = [vp]
    lib.sgz_spectrum_get_raw_fft.argtypes = [vp, fp, i32]
    lib.sgz_spectrum_get_transform_result.argtypes = [vp, fp, i32]
    lib.sgz_spectrum_get_results.argtypes = [vp, i32, fp, i32]
    lib.sgz_spectrum_get_states.argtypes = [vp, i32, fp, i32]
    lib.sgz_spectrum_set_states.argtypes = [vp, i32, fp, i32]
    lib.sgz_spectrum_process_frames.argtypes = [vp, fp, i32, i32, u64, u64, u32, u32, fp, i32]
    lib.sgz_spectrum_process_frames_select.argtypes = [vp, fp, i32, i32, u64, u64, u32, u32, fp, i32, i32]
    lib.sgz_context_create.argtypes = [i32, C.POINTER(vp)]
    lib.sgz_context_destroy.argtypes = [vp]
    lib.sgz_context_destroy.restype = None
    lib.sgz_context_set_cuda_stream.argtypes = [vp, vp]
    lib.sgz_context_synchronize.argtypes = [vp]
    vc = C.POINTER(_abi.VScopeConfig)
    lib.sgz_vscope_vertices.argtypes = [vp, vc, fp, i32, i32, u32, u64, u64, fp, i32]
    lib.sgz_vscope_histogram.argtypes = [vp, vc, fp, i32, i32, u32, u64, u64, fp, i32]
    lib.sgz_vscope_meters.argtypes = [vp, fp, i32, i32, u32, u64, u64, C.c_float, C.c_float, C.c_float, C.POINTER(_abi.VScopeFilters)]
    rc_ = C.POINTER(_abi.ScopeResampleConfig)
    lib.sgz_scope_resample_count.argtypes = [rc_]
    lib.sgz_scope_resample_count.restype = C.c_uint64
    lib.sgz_scope_resample.argtypes = [vp, rc_, C.POINTER(C.c_double), fp, i32, i32, u32, u64, u64, fp, i32]
    lib.sgz_vscope_peak_envelope.argtypes = [vp, fp, i32, i32, u32, u64, u64, C.c_float, C.c_double, C.POINTER(_abi.VScopeFilters), C.POINTER(C.c_double)]
    lib.sgz_scope_trigger_windows.argtypes = [u32, C.POINTER(_abi.ScopeWindowState), C.POINTER(u64), C.POINTER(u32), u32, u64, u64,
                                              C.POINTER(_abi.ScopeWindowEvent), C.POINTER(u32), u32]
    lib.sgz_scope_trigger.argtypes = [vp, i32, i32, fp, i32, i32, u32, u64, u64, C.c_double, C.c_double, u64,
                                      C.POINTER(_abi.ScopeTriggerState), C.POINTER(C.c_uint64), C.POINTER(u32), u32]
    lib.sgz_scope_trigger_device.argtypes = [vp, i32, i32, fp, i32, u32, u64, u64, C.c_double, C.c_double, u64, vp, vp, vp, u32]
    lib.sgz_scope_period.argtypes = [vp, C.POINTER(_abi.ScopePeriodParams), fp, i32, i32, u32, u64, u64,
                                     C.POINTER(_abi.ScopePeriodState), C.POINTER(_abi.ScopePeriodResult)]
    lib.sgz_scope_colour.argtypes = [vp, C.POINTER(_abi.ScopeColourConfig), fp, i32, i32, u32, u64, u64, C.POINTER(_abi.ScopeColourState), fp, i32, C.POINTER(C.c_double)]
    lib.sgz_scope_peak_envelope.argtypes = [vp, i32, fp, i32, i32, u32, u64, u64, C.c_double, C.POINTER(C.c_float), C.POINTER(C.c_double)]
    lib.sgz_spectrum_audio_entry_point.argtypes = [vp, fp, u64, u64, fp, u64, u64, i32, i32, fp, u32, C.POINTER(u32)]
    lib.sgz_spectrum_colour_columns.argtypes = [vp, fp, i32, u32, u32, fp, fp, fp, i32]
    lib.sgz_resonator_process.argtypes = [vp, fp, i32, i32, u64, u64]
    lib.sgz_resonator_get_state.argtypes = [vp, fp, i32]
    lib.sgz_resonator_set_path.argtypes = [vp, i32]
    lib.sgz_spectrum_reserve.argtypes = [vp, u64, u64, u32]
    lib.sgz_fft_forward_noalloc.argtypes = [fp, fp, u32, u32, vp]
    lib.sgz_spectrum_track_peaks.argtypes = [vp, i32, C.POINTER(C.c_double), C.c_double, C.c_double, C.POINTER(_abi.TrackerState), C.POINTER(_abi.TrackerResult)]
    lib.sgz_spectrum_history_reserve.argtypes = [vp, u64]
    lib.sgz_spectrum_history_append.argtypes = [vp, fp, i32, i32, u64, u64]
    lib.sgz_spectrum_history_info.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
    lib.sgz_spectrum_history_view.argtypes = [vp, u64, C.POINTER(vp), C.POINTER(u64)]
    lib.sgz_spectrum_history_read.argtypes = [vp, u64, fp, i32]
    lib.sgz_spectrum_history_process.argtypes = [vp, u32, u32, fp, i32, C.POINTER(u32)]
    for n in EXPORTS:
        f = getattr(lib, n)  # raises AttributeError if the symbol is not exported
        if n not in ("sgz_last_error", "sgz_spectrum_default_config", "sgz_spectrum_destroy", "sgz_context_destroy", "sgz_scope_resample_count",
                     "sgz_spectrum_kernel_name"):
            f.restype = C.c_int
    _lib = lib
    return lib


def check(rc: int, allow=(0,)):
    if rc not in allow:
        raise SgzError(rc, load().sgz_last_error().decode(errors="replace"))
    return rc
