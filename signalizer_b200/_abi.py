"""ctypes mirror of include/sgz.h (enums, PODs).  Keep in lock-step with the header."""
import ctypes as C

ABI_VERSION = 1

# status codes
OK, SKIPPED, EINVAL, ECUDA, ENOMEM, EUNSUPPORTED, ESTATE = 0, 1, -1, -2, -3, -4, -5

# SpectrumChannels — Source/Common/CommonSignalizer.h:495-539
CH_LEFT, CH_RIGHT, CH_MERGE, CH_SIDE, CH_PHASE, CH_SEPARATE, CH_MIDSIDE, CH_COMPLEX = range(8)
CHANNEL_NAMES = ["Left", "Right", "Merge", "Side", "Phase", "Separate", "MidSide", "Complex"]
# SpectrumParameters.h:48-75
BIN_NONE, BIN_LINEAR, BIN_LANCZOS = range(3)
DISPLAY_LINEGRAPH, DISPLAY_COLOURSPECTRUM = range(2)
ALGO_FFT, ALGO_RSNT = range(2)
VIEW_LINEAR, VIEW_LOGARITHMIC = range(2)
LINE_MAIN, LINE_SECOND, LINE_END = range(3)
(WIN_RECTANGULAR, WIN_HANN, WIN_HAMMING, WIN_BLACKMAN, WIN_BLACKMAN_HARRIS, WIN_NUTTALL,
 WIN_FLATTOP, WIN_END) = range(8)
WINSHAPE_SYMMETRIC, WINSHAPE_PERIODIC = range(2)
MEM_HOST, MEM_DEVICE = range(2)
LAYOUT_PLANAR, LAYOUT_INTERLEAVED = range(2)


class SpectrumConfig(C.Structure):
    """sgz_spectrum_config — the fields of TransformConstant<T> the path reads
    (Source/Spectrum/TransformConstant.h:188-239) and the inputs of its table builders."""
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("sample_rate", C.c_float),
        ("window_size", C.c_uint32),
        ("axis_points", C.c_uint32),
        ("algo", C.c_int32),
        ("configuration", C.c_int32),
        ("bin_polation", C.c_int32),
        ("display_mode", C.c_int32),
        ("view_scaling", C.c_int32),
        ("view_left", C.c_double),
        ("view_right", C.c_double),
        ("min_freq", C.c_double),
        ("slope_a", C.c_double),
        ("slope_b", C.c_double),
        ("low_dbs", C.c_double),
        ("high_dbs", C.c_double),
        ("clip_db", C.c_double),
        ("pole", C.c_float * 2),
        ("window_type", C.c_int32),
        ("window_shape", C.c_int32),
        ("sample_buffer_size", C.c_uint32),
        ("free_q", C.c_int32),
    ]


class SpectrumInfo(C.Structure):
    _fields_ = [
        ("transform_size", C.c_uint32),
        ("window_size", C.c_uint32),
        ("axis_points", C.c_uint32),
        ("n_streams", C.c_uint32),
        ("state_channels", C.c_uint32),
        ("break_point", C.c_uint32),
        ("window_kernel_scale", C.c_float),
        ("inv_size", C.c_float),
        ("freq_to_bin", C.c_float),
        ("resonator_vectors", C.c_uint32),
        ("device", C.c_int32),
        ("fused_available", C.c_int32),
    ]


def default_config(**overrides) -> SpectrumConfig:
    """Defaults of SURVEY.md §8(d): Separate, FFT, Lanczos, log view 10 Hz..fs/2, slope a=0 b=1,
    -120..0 dB, clip -384 dB (SpectrumParameters.h:78), poles {0.95, 0.99}."""
    cfg = SpectrumConfig()
    cfg.abi_version = ABI_VERSION
    cfg.sample_rate = 48000.0
    cfg.window_size = 8192
    cfg.axis_points = 1024
    cfg.algo = ALGO_FFT
    cfg.configuration = CH_SEPARATE
    cfg.bin_polation = BIN_LANCZOS
    cfg.display_mode = DISPLAY_LINEGRAPH
    cfg.view_scaling = VIEW_LOGARITHMIC
    cfg.view_left, cfg.view_right = 0.0, 1.0
    cfg.min_freq = 10.0
    cfg.slope_a, cfg.slope_b = 0.0, 1.0
    cfg.low_dbs, cfg.high_dbs, cfg.clip_db = -120.0, 0.0, -384.0
    cfg.pole[0], cfg.pole[1] = 0.95, 0.99
    cfg.window_type = WIN_BLACKMAN_HARRIS
    cfg.window_shape = WINSHAPE_PERIODIC
    cfg.sample_buffer_size = 200
    cfg.free_q = 0
    for k, v in overrides.items():
        if k == "pole":
            cfg.pole[0], cfg.pole[1] = v
        else:
            if not hasattr(cfg, k):
                raise AttributeError(k)
            setattr(cfg, k, v)
    return cfg


def copy_config(cfg: SpectrumConfig, **overrides) -> SpectrumConfig:
    out = SpectrumConfig()
    C.memmove(C.byref(out), C.byref(cfg), C.sizeof(SpectrumConfig))
    for k, v in overrides.items():
        if k == "pole":
            out.pole[0], out.pole[1] = v
        else:
            setattr(out, k, v)
    return out


# ---- secondary paths (include/sgz.h: Vectorscope / Oscilloscope) --------------------------------
OSC_LEFT, OSC_RIGHT, OSC_MID, OSC_SIDE = range(4)


class VScopeConfig(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("polar", C.c_int32), ("scale_polar", C.c_int32), ("gain", C.c_float),
                ("width", C.c_uint32), ("height", C.c_uint32)]


class VScopeFilters(C.Structure):
    _fields_ = [("envelope", C.c_float * 2), ("balance", (C.c_float * 2) * 2), ("phase", C.c_float * 2)]


class ScopeResampleConfig(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("channel", C.c_int32), ("samples_per_pixel", C.c_double),
                ("left", C.c_double), ("right", C.c_double), ("inc", C.c_double)]


class ScopeTriggerState(C.Structure):
    _fields_ = [("state", C.c_double), ("cross_origin", C.c_uint64), ("is_peak_holding", C.c_int32), ("reserved", C.c_int32)]


class ScopeWindowState(C.Structure):
    _fields_ = [("window_size", C.c_double), ("old_peak", C.c_uint64), ("current_peak", C.c_uint64), ("buffered_samples", C.c_uint64),
                ("front_origin", C.c_uint64), ("steady_clock", C.c_uint64), ("is_working_on_peak", C.c_int32), ("window_changed", C.c_int32)]


class ScopeWindowEvent(C.Structure):
    _fields_ = [("origin", C.c_uint64), ("size", C.c_uint64), ("buffered", C.c_uint64), ("kind", C.c_uint32), ("reserved", C.c_uint32)]


WINDOW_PROCESS, WINDOW_SWAP = 0, 1
OUT_FULL, OUT_LINE_MAIN, OUT_LAST_FRAME = 0, 1, 2   # sgz_spectrum_process_frames_select


class ScopeBinRecord(C.Structure):
    _fields_ = [("index", C.c_uint64), ("value", C.c_double), ("offset", C.c_double)]


class ScopePeriodState(C.Structure):
    _fields_ = [("median", ScopeBinRecord * 8), ("median_pos", C.c_uint64)]


class ScopePeriodParams(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("channel", C.c_int32), ("effective_window_size", C.c_double), ("sample_rate", C.c_double),
                ("threshold", C.c_double), ("hysteresis", C.c_double), ("phase_offset_degrees", C.c_double)]


class ScopePeriodResult(C.Structure):
    _fields_ = [("record", ScopeBinRecord), ("fundamental", C.c_double), ("cycle_samples", C.c_double), ("phase", C.c_double),
                ("sample_offset", C.c_double)]


def vscope_config(polar=True, scale_polar=False, gain=1.0, width=512, height=512) -> VScopeConfig:
    return VScopeConfig(ABI_VERSION, int(polar), int(scale_polar), gain, width, height)


def resample_config(channel=OSC_LEFT, samples_per_pixel=1.0 / 16, left=0.0, right=1.0, inc=1.0 / 1024) -> ScopeResampleConfig:
    return ScopeResampleConfig(ABI_VERSION, channel, samples_per_pixel, left, right, inc)


# frequency tracker (include/sgz.h: sgz_tracker_result / sgz_tracker_state)
TRACK_LINE_MAIN, TRACK_LINE_SECOND, TRACK_TRANSFORM = range(3)


class TrackerResult(C.Structure):
    _fields_ = [("peak_offset", C.c_int64), ("searched_raw_fft", C.c_int32), ("peak_is_complex", C.c_int32),
                ("peak_frequency", C.c_double), ("peak_dbs", C.c_double), ("peak_deviance", C.c_double), ("peak_slope", C.c_double),
                ("peak_x", C.c_double), ("peak_fraction", C.c_double), ("smoothed_frequency", C.c_double), ("smoothed_dbs", C.c_double)]


class TrackerState(C.Structure):
    _fields_ = [("frequency_state", C.c_double * 8), ("peak_dbs_state", C.c_double * 8),
                ("current_frequency", C.c_double), ("current_peak_dbs", C.c_double), ("current_peak", C.c_double)]


# oscilloscope colouring / envelopes (include/sgz.h: sgz_scope_colour_config / _state)
OSC_SEPARATE, OSC_MIDSIDE = 4, 5
ENVELOPE_NONE, ENVELOPE_RMS, ENVELOPE_PEAK = range(3)


class ScopeColourConfig(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("channel_mode", C.c_int32), ("envelope_mode", C.c_int32), ("sample_rate", C.c_double),
                ("colour_smoothing_ms", C.c_double), ("envelope_window", C.c_double), ("colouring_blend", C.c_double),
                ("colours", (C.c_float * 3) * 3), ("key_left", C.c_uint8 * 4), ("key_right", C.c_uint8 * 4)]


class ScopeColourState(C.Structure):
    _fields_ = [("smooth_filters", C.c_float * 3), ("aux_smooth_filter", C.c_float * 3), ("network", C.c_float * 24), ("envelope", C.c_float)]


def colour_config(channel_mode=OSC_LEFT, envelope_mode=ENVELOPE_RMS, sample_rate=48000.0, colour_smoothing_ms=1.0, envelope_window=0.1,
                  colouring_blend=1.0, colours=((1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0)), key_left=(0, 255, 0, 255), key_right=(255, 0, 0, 255)):
    c = ScopeColourConfig()
    c.abi_version = ABI_VERSION
    c.channel_mode, c.envelope_mode, c.sample_rate = channel_mode, envelope_mode, sample_rate
    c.colour_smoothing_ms, c.envelope_window, c.colouring_blend = colour_smoothing_ms, envelope_window, colouring_blend
    for i in range(3):
        for k in range(3):
            c.colours[i][k] = colours[i][k]
    for k in range(4):
        c.key_left[k], c.key_right[k] = key_left[k], key_right[k]
    return c
