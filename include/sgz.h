/*
 * sgz.h — C ABI of the B200-native Signalizer analysis hot path.
 *
 * This is the drop-in boundary for the Spectrum transform pipeline of
 * jthorborg/signalizer (reference paths are relative to the reference
 * checkout, `Source/...`).  The reference exposes the path as the C++
 * template pair TransformConstant<T> / TransformPair<T>; there is no C ABI
 * upstream, so every entry point below names the C++ member it replaces.
 *
 * All entry points are batched over the `n_streams` stereo pairs of a handle
 * (the reference loops `parallel_for(numChannels / 2, ...)` over at most 8
 * pairs, SpectrumDSP.cpp:83-102).  A handle owns all device memory; the caller
 * owns host buffers (and device buffers it passes with SGZ_MEM_DEVICE).  A
 * handle is externally synchronised like the reference's `streamState` mutex
 * (SpectrumDSP.cpp:67, SpectrumRendering.cpp:594): one caller at a time.
 *
 * Nothing here has a CPU fallback: without a CUDA device every compute entry
 * point returns SGZ_ECUDA.
 */
#ifndef SGZ_H
#define SGZ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SGZ_ABI_VERSION 1

/* ---- status codes -------------------------------------------------------
 * The reference signals "skip this frame" with `return false`
 * (TransformDSP.inl:45-46, :241-242), silently returns on degenerate sizes
 * (:519-520) and throws on invariants (:1301, :613-616).  Nothing is thrown
 * across this ABI. */
enum {
    SGZ_OK           = 0,
    SGZ_SKIPPED      = 1,  /* prepareTransform() == false */
    SGZ_EINVAL       = -1,
    SGZ_ECUDA        = -2,
    SGZ_ENOMEM       = -3,
    SGZ_EUNSUPPORTED = -4,
    SGZ_ESTATE       = -5  /* call order violated (TransformPair.h:123-124) */
};

/* SpectrumChannels, Source/Common/CommonSignalizer.h:495-539 (same values). */
enum {
    SGZ_CH_LEFT = 0, SGZ_CH_RIGHT = 1, SGZ_CH_MERGE = 2, SGZ_CH_SIDE = 3,
    SGZ_CH_PHASE = 4, SGZ_CH_SEPARATE = 5, SGZ_CH_MIDSIDE = 6, SGZ_CH_COMPLEX = 7
};
/* SpectrumContent::BinInterpolation, Source/Spectrum/SpectrumParameters.h:53-58 */
enum { SGZ_BIN_NONE = 0, SGZ_BIN_LINEAR = 1, SGZ_BIN_LANCZOS = 2 };
/* SpectrumContent::DisplayMode, SpectrumParameters.h:60-64 */
enum { SGZ_DISPLAY_LINEGRAPH = 0, SGZ_DISPLAY_COLOURSPECTRUM = 1 };
/* SpectrumContent::TransformAlgorithm, SpectrumParameters.h:66-69 */
enum { SGZ_ALGO_FFT = 0, SGZ_ALGO_RSNT = 1 };
/* SpectrumContent::ViewScaling, SpectrumParameters.h:71-75 */
enum { SGZ_VIEW_LINEAR = 0, SGZ_VIEW_LOGARITHMIC = 1 };
/* SpectrumContent::LineGraphs, SpectrumParameters.h:48-51 */
enum { SGZ_LINE_MAIN = 0, SGZ_LINE_SECOND = 1, SGZ_LINE_END = 2 };

/* cpl::dsp::WindowTypes (cpl is absent from the reference checkout; the set
 * below is the cosine-sum family the restatement defines). */
enum {
    SGZ_WIN_RECTANGULAR = 0, SGZ_WIN_HANN = 1, SGZ_WIN_HAMMING = 2,
    SGZ_WIN_BLACKMAN = 3, SGZ_WIN_BLACKMAN_HARRIS = 4, SGZ_WIN_NUTTALL = 5,
    SGZ_WIN_FLATTOP = 6, SGZ_WIN_END = 7
};
/* cpl::dsp::Windows::Shape */
enum { SGZ_WINSHAPE_SYMMETRIC = 0, SGZ_WINSHAPE_PERIODIC = 1 };

/* `device` value for sgz_spectrum_create that builds the host-side tables only (for inspecting the
 * derived constants without a GPU).  Every compute entry point of such a handle returns SGZ_ECUDA. */
#define SGZ_DEVICE_NONE (-1)

/* where a caller-provided buffer lives */
enum { SGZ_MEM_HOST = 0, SGZ_MEM_DEVICE = 1 };
/* audio layouts: PLANAR is the reference's `float** buffer`, pair i =
 * channels 2i, 2i+1 (SpectrumDSP.cpp:98) flattened to [stream][2][sample];
 * INTERLEAVED is [stream][sample][2]. */
enum { SGZ_LAYOUT_PLANAR = 0, SGZ_LAYOUT_INTERLEAVED = 1 };

/* Everything TransformConstant<T> holds that the hot path reads
 * (Source/Spectrum/TransformConstant.h:188-239) plus the inputs of the
 * functions that derive its tables (setStorage :81-92, remapFrequencies
 * :125-180, generateSlopeMap :109-118, regenerateWindowKernel :104-107,
 * remapResonator :120-123; driven from Spectrum.cpp:365-394,489,571-595). */
typedef struct sgz_spectrum_config {
    uint32_t abi_version;        /* SGZ_ABI_VERSION */
    float    sample_rate;        /* Constant::sampleRate (float, Spectrum.cpp:430-431) */
    uint32_t window_size;        /* effectiveWindowSize; transformSize = max(32, nextPow2(window)) */
    uint32_t axis_points;        /* elements == display points == filters */
    int32_t  algo;               /* SGZ_ALGO_* */
    int32_t  configuration;      /* SGZ_CH_* */
    int32_t  bin_polation;       /* SGZ_BIN_* */
    int32_t  display_mode;       /* SGZ_DISPLAY_* */
    int32_t  view_scaling;       /* SGZ_VIEW_* */
    double   view_left;          /* viewRect.left  (0..1) */
    double   view_right;         /* viewRect.right (0..1) */
    double   min_freq;           /* state.minLogFreq, 10 Hz (Spectrum.cpp:83) */
    double   slope_a;            /* PowerFunction{a,b}: slope = b * f^a */
    double   slope_b;
    double   low_dbs;            /* Constant::lowDBs */
    double   high_dbs;           /* Constant::highDBs */
    double   clip_db;            /* Constant::clipDB, kMinDbs = -384 */
    float    pole[2];            /* Constant::filter[k].pole, k = LineMain, LineSecond */
    int32_t  window_type;        /* SGZ_WIN_* */
    int32_t  window_shape;       /* SGZ_WINSHAPE_* */
    uint32_t sample_buffer_size; /* Constant::sampleBufferSize ("blob"), default 200 */
    int32_t  free_q;             /* remapResonator(shouldHaveFreeQ, ...) */
} sgz_spectrum_config;

typedef struct sgz_spectrum_info {
    uint32_t transform_size;     /* N */
    uint32_t window_size;        /* W */
    uint32_t axis_points;        /* P */
    uint32_t n_streams;
    uint32_t state_channels;     /* getStateConfigurationChannels(): 1 or 2 */
    uint32_t break_point;        /* first display point handled by max-of-bins */
    float    window_kernel_scale;
    float    inv_size;           /* windowKernelScale / (W * 0.5) */
    float    freq_to_bin;
    uint32_t resonator_vectors;  /* windowCoefficients(window).second */
    int32_t  device;
    int32_t  fused_available;    /* 1 if sgz_spectrum_process_frames has a fused kernel for this config */
} sgz_spectrum_info;

typedef struct sgz_spectrum sgz_spectrum;

/* Fills `cfg` with the defaults of SURVEY.md §8(d) (Separate, FFT, Lanczos,
 * log view, -120..0 dB, poles {0.95, 0.99}). */
void sgz_spectrum_default_config(sgz_spectrum_config* cfg);

/* Replaces: `TransformConstant<T>` construction + `handleFlagUpdates`
 * (Spectrum.cpp:351-616) + `pairs.resize(channels / 2)`
 * (SpectrumDSP.cpp:225).  Builds every table on the host and uploads it. */
int  sgz_spectrum_create(const sgz_spectrum_config* cfg, uint32_t n_streams, int device,
                         sgz_spectrum** out);
void sgz_spectrum_destroy(sgz_spectrum* h);
int  sgz_spectrum_get_info(const sgz_spectrum* h, sgz_spectrum_info* out);
const char* sgz_last_error(void);

/* The fused kernel sgz_spectrum_process_frames launches for this handle, as a
 * profiler prints it ("sgz::k_fused8<8, 1, 2, 0, 0>"; "" when the configuration
 * runs the staged kernels).  No reference counterpart: measurement only. */
const char* sgz_spectrum_kernel_name(const sgz_spectrum* h);
/* All work of a handle is enqueued on one CUDA stream (default: a stream the
 * handle creates).  `cuda_stream` is a cudaStream_t. */
int  sgz_spectrum_set_cuda_stream(sgz_spectrum* h, void* cuda_stream);
int  sgz_spectrum_synchronize(sgz_spectrum* h);

/* Host copies of the derived tables (bit-exact parity targets).
 * windowKernel[N] (TransformConstant.h:220), mappedFrequencies[P] (:215),
 * slopeMap[P] (:232). */
int  sgz_spectrum_get_window_kernel(const sgz_spectrum* h, float* out, size_t n);
int  sgz_spectrum_get_mapped_frequencies(const sgz_spectrum* h, float* out, size_t n);
int  sgz_spectrum_get_slope_map(const sgz_spectrum* h, float* out, size_t n);
/* Per display point: mode[x] (0 = interpolate, 1 = max-of-bins), and the
 * integer bin ranges the loops of TransformDSP.inl:568-639 visit:
 * first[x] / count[x] = bins scanned (max) or nearest/floor index
 * (interpolate), fallback[x] = the `maxLBin = bin` initial value. */
int  sgz_spectrum_get_bin_map(const sgz_spectrum* h, int32_t* mode, int32_t* first,
                              int32_t* count, int32_t* fallback, size_t n);

/* ---- staged pipeline: TransformPair.h:123-157 ---------------------------
 * `audio` holds, per stream, `n_samples` samples per channel of history,
 * oldest first; the last window_size samples are used
 * (TransformDSP.inl:49).  stream_stride is in floats between streams.
 * Replaces `bool prepareTransform(const Constant&, const AudioPair&)`
 * (TransformDSP.inl:39-231).  Returns SGZ_SKIPPED when n_samples <
 * window_size. */
int  sgz_spectrum_prepare(sgz_spectrum* h, const float* audio, int mem, int layout,
                          uint64_t n_samples, uint64_t stream_stride);
/* Replaces the preliminary-audio overload (TransformDSP.inl:234-484): the
 * timeline is `history` (n_history samples/channel) followed by `prelim`
 * (n_prelim samples/channel, newest). */
int  sgz_spectrum_prepare_preliminary(sgz_spectrum* h, const float* history, uint64_t n_history,
                                      uint64_t history_stride, const float* prelim,
                                      uint64_t n_prelim, uint64_t prelim_stride, int mem,
                                      int layout);
/* Replaces `void doTransform(const Constant&)` (TransformDSP.inl:487-502). */
int  sgz_spectrum_transform(sgz_spectrum* h);
/* Replaces `template<ISA> void mapToLinearSpace(const Constant&)`
 * (TransformDSP.inl:506-1135). */
int  sgz_spectrum_map(sgz_spectrum* h);
/* Replaces `void postProcessStdTransform(const Constant&)`
 * (TransformDSP.inl:1438-1441 -> mapAndTransformDFTFilters :1299-1435). */
int  sgz_spectrum_postprocess(sgz_spectrum* h);

/* Replaces `getRawFFT` (TransformPair.h:164-167): [stream][N] complex. */
int  sgz_spectrum_get_raw_fft(sgz_spectrum* h, float* out, int mem);
/* Replaces `getTransformResult` (TransformDSP.inl:1444-1447):
 * [stream][P * state_channels * 2] floats. */
int  sgz_spectrum_get_transform_result(sgz_spectrum* h, float* out, int mem);
/* Replaces `lineGraphs[graph].getResults(size)` (TransformPair.h:72-76):
 * [stream][P] UComplexFilter<float> = 2 floats each
 * (CommonSignalizer.h:541-591). */
int  sgz_spectrum_get_results(sgz_spectrum* h, int graph, float* out, int mem);
/* The peak-hold states behind the results (TransformPair.h:92). */
int  sgz_spectrum_get_states(sgz_spectrum* h, int graph, float* out, int mem);
int  sgz_spectrum_set_states(sgz_spectrum* h, int graph, const float* in, int mem);
/* Replaces `clearLineGraphStates` / `clearAudioState` (TransformPair.h:169-184). */
int  sgz_spectrum_clear_line_graph_states(sgz_spectrum* h);
int  sgz_spectrum_clear_audio_state(sgz_spectrum* h);

/* Sizes, once, every buffer the process calls below would otherwise grow on first use: the device staging of
 * host-resident audio (`n_samples` per channel and stream at `stream_stride`) and results (`n_frames` frames, full
 * output), the copy streams and events of the host-buffer pipeline, the resonator's time-slice scratch.  After it,
 * sgz_spectrum_process_frames* / sgz_resonator_process / sgz_spectrum_audio_entry_point with sizes up to these call
 * neither cudaMalloc nor cudaFree (the reference's TransformPair likewise grows its vectors lazily,
 * TransformPair.h:213-233; a caller with a deadline reserves up front).  Optional: without it the first call, and
 * any call with larger sizes, allocates and may wait for the handle's stream. */
int  sgz_spectrum_reserve(sgz_spectrum* h, uint64_t n_samples, uint64_t stream_stride, uint32_t n_frames);

/* ---- fused batched path --------------------------------------------------
 * One call = for every stream, `n_frames` successive frames: frame f uses
 * samples [f*hop, f*hop + window_size) of the stream and runs
 * prepareTransform -> doTransform -> mapToLinearSpace ->
 * postProcessStdTransform (the body of the loop at
 * SpectrumRendering.cpp:622-634, and of `addAudioFrame`,
 * TransformDSP.inl:1139-1148, at every `sampleBufferSize` boundary of
 * audioEntryPoint :1172-1204 when hop == sample_buffer_size).
 * `results`: [stream][frame][graph(2)][P][2] floats.  Peak-hold state is
 * carried frame to frame and call to call.  n_samples must be >=
 * (n_frames-1)*hop + window_size.
 * One fused kernel launch covers all frames of all streams when the
 * configuration has a fused kernel (sgz_spectrum_info.fused_available:
 * FFT algorithm, transform size 1024 ... 16384, display points that fit one
 * CTA's shared memory): the launch is as many CTAs as the GPU holds at once,
 * each taking an equal share of the (stream, frame) list, whatever the stream
 * count, with bit-identical results (the peak-hold state is handed across
 * where a stream is cut between CTAs).  Everything else runs the four staged
 * kernels per frame. */
int  sgz_spectrum_process_frames(sgz_spectrum* h, const float* audio, int mem, int layout,
                                 uint64_t n_samples, uint64_t stream_stride, uint32_t hop,
                                 uint32_t n_frames, float* results, int results_mem);
/* The same call with a choice of what is written to `results` (the transfer
 * back to a host caller is most of the cost of the call above):
 *   SGZ_OUT_FULL        [stream][frame][graph(2)][P][2]      (= sgz_spectrum_process_frames)
 *   SGZ_OUT_LINE_MAIN   [stream][frame][P][2]: graph LineMain of every frame, what
 *                       `addAudioFrame` appends to `sfbuf` (TransformDSP.inl:1139-1148) and the
 *                       spectrogram consumes (SpectrumDSP.cpp:119-199)
 *   SGZ_OUT_LAST_FRAME  [stream][graph(2)][P][2]: the last frame only, what the line graph shows
 *                       (`lineGraphs[k].getResults`, TransformPair.h:72-76)
 * Every frame is computed in all three; states and sgz_spectrum_get_results are the same after the call. */
enum { SGZ_OUT_FULL = 0, SGZ_OUT_LINE_MAIN = 1, SGZ_OUT_LAST_FRAME = 2 };
int  sgz_spectrum_process_frames_select(sgz_spectrum* h, const float* audio, int mem, int layout,
                                        uint64_t n_samples, uint64_t stream_stride, uint32_t hop,
                                        uint32_t n_frames, float* results, int results_mem, int output);

/* ---- audioEntryPoint ------------------------------------------------------
 * Replaces `template<ISA> void audioEntryPoint(const Constant&, const
 * optional<AudioPair>&, array<AFloat*, 2> buffer, size_t numSamples)`
 * (TransformDSP.inl:1165-1211) as called by AudioDispatcher::dispatch
 * (SpectrumDSP.cpp:81-100) for one callback block, statement by statement:
 * in ColourSpectrum display mode the block is cut at `sample_buffer_size`
 * boundaries (the counter `processedSamplesSinceLastFrame` is carried in the
 * handle); RSNT resonates each piece; at every boundary FFT runs
 * prepareTransform(history, buffer, availableSamples) + doTransform, then
 * addAudioFrame (:1139-1148) appends LineMain's results.  In LineGraph mode
 * only RSNT resonates the block.  `history` is the audio-history view (FFT
 * only; may be NULL for RSNT), `block` the callback buffer, both laid out
 * like sgz_spectrum_prepare.  frames: HOST [stream][max_frames][P][2];
 * *n_emitted receives the number of frames produced (the sfbuf size). */
int  sgz_spectrum_audio_entry_point(sgz_spectrum* h, const float* history, uint64_t n_history,
                                    uint64_t history_stride, const float* block, uint64_t n_block,
                                    uint64_t block_stride, int mem, int layout, float* frames,
                                    uint32_t max_frames, uint32_t* n_emitted);

/* ---- device-resident audio history --------------------
 * Replaces the audio-history side of the path: the reference reads the last
 * `windowSize` samples of a `cpl::AudioStream` history through its two-segment
 * views (`audio[c].getItIndex(k)` over two ranges, TransformDSP.inl:49,65-88;
 * the ring wraps, hence two segments) and is called back per block
 * (audioEntryPoint, :1165-1211).  Here the history is a per-stream ring in
 * HBM: only new samples are uploaded, windows are never re-sent, and the
 * fused frame kernel reads the ring in place.
 *   reserve : allocate `capacity` (>= window_size) samples per stream,
 *             cleared to silence; resets the counters.  Calling it again
 *             clears the history.
 *   append  : add one callback block ([stream][n_block] stereo, HOST or
 *             DEVICE, either layout) at the write head, wrapping as needed.
 *   view    : device pointer to the last n_last (<= capacity) samples of
 *             stream 0 as ONE contiguous interleaved run (the two segments
 *             linearised), `*stream_stride` floats between streams; valid
 *             until the next append; usable with every entry point that
 *             takes SGZ_MEM_DEVICE audio (prepare, audio_entry_point ...).
 *   read    : copy of that run, [stream][n_last][2] floats.
 *   process : every frame that fell due since the last call — one per `hop`
 *             appended samples, the frame at boundary b transforming samples
 *             [b - window_size, b) (silence before time zero), exactly the
 *             frames audioEntryPoint emits when sample_buffer_size == hop —
 *             runs through the fused path of sgz_spectrum_process_frames;
 *             at most max_frames per call (the rest stay pending);
 *             results as there, [stream][*n_frames][graph][P][2].  SGZ_ESTATE
 *             if a pending frame's window has already been overwritten. */
int  sgz_spectrum_history_reserve(sgz_spectrum* h, uint64_t capacity);
int  sgz_spectrum_history_append(sgz_spectrum* h, const float* block, int mem, int layout,
                                 uint64_t n_block, uint64_t stream_stride);
int  sgz_spectrum_history_info(sgz_spectrum* h, uint64_t* capacity, uint64_t* total_appended,
                               uint64_t* last_frame_position);
int  sgz_spectrum_history_view(sgz_spectrum* h, uint64_t n_last, const float** device_audio,
                               uint64_t* stream_stride);
int  sgz_spectrum_history_read(sgz_spectrum* h, uint64_t n_last, float* out, int mem);
int  sgz_spectrum_history_process(sgz_spectrum* h, uint32_t hop, uint32_t max_frames,
                                  float* results, int results_mem, uint32_t* n_frames);

/* ---- frequency tracker --------------------------------
 * Replaces the peak search of `Spectrum::drawFrequencyTracking`
 * (Source/Spectrum/SpectrumRendering.cpp:301-469, 494-519) and
 * `SmoothedPeakState` (Source/Spectrum/Spectrum.h:405-459) for every stream at
 * once; the mouse read-out, the scalloping-loss estimate and the text box
 * around it are GUI and stay with the caller.
 * `graph` = state.frequencyTrackingGraph.  SGZ_TRACK_TRANSFORM with the FFT
 * algorithm and a non-Complex configuration searches the raw FFT
 * (`transform.getRawFFT`, :376-469: |z|^2 argmax in the bins around the
 * mouse, rising-edge scan at the window borders, log-domain parabolic fit);
 * that needs the handle's audioMemory as the staged calls leave it, i.e.
 * after sgz_spectrum_transform / _map.  Every other case searches
 * lineGraphs[graph].getResults() (:306-374; TRANSFORM falls back to
 * LINE_MAIN).  mouse_fraction: HOST [stream] = mouseX / (width - 1).
 * states: HOST [stream], in/out, NULL = no smoothing (the smoothed_* fields
 * then repeat the raw values); smoothing_ms = trackerSmoothing * 1000, fps as
 * passed to drawFrequencyTracking.  results: HOST [stream]. */
enum { SGZ_TRACK_LINE_MAIN = 0, SGZ_TRACK_LINE_SECOND = 1, SGZ_TRACK_TRANSFORM = 2 };
typedef struct sgz_tracker_result {
    int64_t peak_offset;        /* index into the searched array */
    int32_t searched_raw_fft;   /* 1 = the raw-FFT branch ran */
    int32_t peak_is_complex;    /* peakFrequency was above Nyquist and got folded (:494-496) */
    double  peak_frequency, peak_dbs, peak_deviance, peak_slope, peak_x, peak_fraction;
    double  smoothed_frequency, smoothed_dbs; /* peakState.getFrequency() / getPeakDBs() */
} sgz_tracker_result;
typedef struct sgz_tracker_state {  /* SmoothedPeakState's members (Spectrum.h:452-455) */
    double frequency_state[8], peak_dbs_state[8];
    double current_frequency, current_peak_dbs, current_peak;
} sgz_tracker_state;
int  sgz_spectrum_track_peaks(sgz_spectrum* h, int graph, const double* mouse_fraction,
                              double smoothing_ms, double fps, sgz_tracker_state* states,
                              sgz_tracker_result* results);

/* ---- spectrogram colour columns -----------------------------------------
 * Replaces `AudioDispatcher::blendAndDispatchSpectrums`
 * (Source/Spectrum/SpectrumDSP.cpp:110-206): every frame of `results`
 * ([stream][frame][graph][P][2], as written by sgz_spectrum_process_frames;
 * LineMain's .magnitude is the intensity) becomes one RGBA8 column of P
 * pixels.  `pairs_per_view` consecutive streams are the stereo pairs of one
 * view and are screen-blended (c += (1 - c) * colour, :162-166) into the same
 * column; n_streams must be a multiple of it.  `ratios` is
 * Constant::normalizedSpecRatios (6 floats, [0] = 0, Spectrum.cpp:226-246),
 * `colours` the per-pair SpectrumColourArray [pairs_per_view][6][3] (background
 * + 5 gradient colours, RGB floats).  rgba: [n_streams / pairs_per_view][frame][P][4]. */
int  sgz_spectrum_colour_columns(sgz_spectrum* h, const float* results, int mem, uint32_t n_frames,
                                 uint32_t pairs_per_view, const float* ratios, const float* colours,
                                 uint8_t* rgba, int out_mem);

/* ---- resonator bank (algo = SGZ_ALGO_RSNT) ------------------------------
 * Replaces `template<ISA> void resonatingDispatch(const Constant&,
 * std::array<AFloat*, 2> buffer, size_t numSamples)` (TransformDSP.inl:1215-1295):
 * advances every stream's complex one-pole bank (filters x window vectors x
 * channels) over a block of `n_samples` samples.  `audio` as in
 * sgz_spectrum_prepare.  The windowed snapshot (copyResonatorStateInto,
 * TransformDSP.inl:1152-1161) is taken by sgz_spectrum_map, and
 * sgz_spectrum_process_frames resonates `hop` samples per frame like
 * audioEntryPoint (:1172-1204). */
int  sgz_resonator_process(sgz_spectrum* h, const float* audio, int mem, int layout,
                           uint64_t n_samples, uint64_t stream_stride);
/* Which kernels advance the resonator bank.  SGZ_RESONATOR_AUTO (default): blocks of at least 2048 samples run on the
 * tensor cores - tiles of 4096 samples as a TF32 GEMM with a three-term operand split and FP32 accumulation - and
 * shorter blocks on the FP32 recurrence; SGZ_RESONATOR_FP32: the FP32 recurrence
 * always (after 480 000 samples the two differ by about 2e-5 of the state's peak - both inside the tolerance held
 * against the CPU oracle - not bit for bit).  No reference counterpart. */
enum { SGZ_RESONATOR_AUTO = 0, SGZ_RESONATOR_FP32 = 1 };
int  sgz_resonator_set_path(sgz_spectrum* h, int path);
/* Raw resonator state (cresonator), [stream][channel][vector][filter] complex. */
int  sgz_resonator_get_state(sgz_spectrum* h, float* out, int mem);

/* ==== secondary paths: Vectorscope and Oscilloscope =========================
 * These views have no per-view constant object; a light context owns the
 * device, the CUDA stream and staging buffers for host-resident callers. */
typedef struct sgz_context sgz_context;
int  sgz_context_create(int device, sgz_context** out);
void sgz_context_destroy(sgz_context* ctx);
int  sgz_context_set_cuda_stream(sgz_context* ctx, void* cuda_stream);
int  sgz_context_synchronize(sgz_context* ctx);

/* channel views of Source/Oscilloscope/SampleColourEvaluators.h:57-320 */
enum { SGZ_OSC_LEFT = 0, SGZ_OSC_RIGHT = 1, SGZ_OSC_MID = 2, SGZ_OSC_SIDE = 3,
       SGZ_OSC_SEPARATE = 4, SGZ_OSC_MIDSIDE = 5 /* the two-trace modes: envelopes only (sgz_scope_colour, sgz_scope_peak_envelope) */ };

/* ---- Vectorscope ----------------------------------------------------------
 * Replaces the per-sample vertex arithmetic of `VectorScope::drawPolarPlot`
 * (Source/Vectorscope/VectorscopeRendering.cpp:500-746) and `drawRectPlot`
 * (:444-497): polar = 135 degree rotation, atan fold, sincos, radius
 * max(|L|, |R|); Lissajous = (R, L); then gain (:512-517) and the optional
 * 0..1 -> -1..1 height transform (:286-300).  xyz: [stream][sample][3]. */
typedef struct sgz_vscope_config {
    uint32_t abi_version;
    int32_t  polar;        /* state.isPolar */
    int32_t  scale_polar;  /* state.scalePolar */
    float    gain;         /* processor->envelopeGain * state.userGain */
    uint32_t width;        /* histogram size (the GL viewport it stands in for) */
    uint32_t height;
} sgz_vscope_config;
int  sgz_vscope_vertices(sgz_context* ctx, const sgz_vscope_config* cfg, const float* audio, int mem,
                         int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
                         float* xyz, int out_mem);
/* The GL_POINTS rasterisation of those vertices (VectorscopeRendering.cpp:547,
 * the reference hands it to OpenGL) as a point-density histogram:
 * hist[stream][row][col] += 1 at floor((c + 1) * size / 2); points outside
 * [-1, 1) are clipped.  `hist` is overwritten.  [stream][height][width]. */
int  sgz_vscope_histogram(sgz_context* ctx, const sgz_vscope_config* cfg, const float* audio, int mem,
                          int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
                          uint32_t* hist, int out_mem);
/* Replaces `VectorScope::Processor::audioProcessing` (Source/Vectorscope/
 * Vectorscope.cpp:268-377): RMS envelope, slow/fast balance and phase
 * correlation one-pole filters.  `filters` (HOST, [stream]) is read, advanced
 * over the block and written back. */
typedef struct sgz_vscope_filters { float envelope[2]; float balance[2][2]; float phase[2]; } sgz_vscope_filters;
int  sgz_vscope_meters(sgz_context* ctx, const float* audio, int mem, int layout, uint32_t n_streams,
                       uint64_t n_samples, uint64_t stream_stride, float envelope_coeff,
                       float stereo_coeff, float second_stereo_filter_speed, sgz_vscope_filters* filters);

/* Replaces `VectorScope::runPeakFilter` (VectorscopeRendering.cpp:826-889): the
 * auto-gain envelope of the vectorscope.  Per stream: the largest |L| and |R|
 * of the window - whole groups of 8 samples only, like the reference's AVX
 * loop (:857-868, its remainder is a TODO at :870) - squared and folded into
 * filters[s].envelope[0 / 1] with coeff = envelope_coeff^(n_samples *
 * delta_time) (:840-842; delta_time = openGLDeltaTime(), seconds per rendered
 * frame), then envelope_gain[s] = 1 / max(sqrt(e0), sqrt(e1)), written only
 * when it is a normal number (:882-888; otherwise the previous gain stays).
 * The result is what sgz_vscope_config.gain takes in the peak envelope mode
 * (gain = envelopeGain * userGain, VectorscopeRendering.cpp:451,512).  filters, envelope_gain: HOST, in/out. */
int  sgz_vscope_peak_envelope(sgz_context* ctx, const float* audio, int mem, int layout, uint32_t n_streams,
                              uint64_t n_samples, uint64_t stream_stride, float envelope_coeff, double delta_time,
                              sgz_vscope_filters* filters, double* envelope_gain);

/* ---- Oscilloscope ---------------------------------------------------------
 * Replaces the Lanczos branch of `Oscilloscope::drawWavePlot`
 * (Source/Oscilloscope/OscilloscopeRendering.cpp:790-891): windowed-sinc
 * (a = 10, 21-tap sliding buffer) resampling of the newest samples to one
 * vertex per output step.  `sample_pos` (HOST, [stream]) is `samplePos` after
 * :826; the buffer's last sample is the newest.  out_xy: [stream][count][2],
 * count = sgz_scope_resample_count(cfg). */
typedef struct sgz_scope_resample_config {
    uint32_t abi_version;
    int32_t  channel;            /* SGZ_OSC_* */
    double   samples_per_pixel;  /* 1 / pixelsPerSample (:824) */
    double   left, right, inc;   /* unit-space range and step (:562-563, :822) */
} sgz_scope_resample_config;
uint64_t sgz_scope_resample_count(const sgz_scope_resample_config* cfg);
int  sgz_scope_resample(sgz_context* ctx, const sgz_scope_resample_config* cfg, const double* sample_pos,
                        const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
                        uint64_t stream_stride, float* out_xy, int out_mem);
/* Replaces `ZeroCrossingProcessor::process` (mode 0) and
 * `PeakHoldProcessor::process` (mode 1) driven by executeSamplingWindows
 * (Source/Oscilloscope/StreamPreprocessing.h:284-347, OscilloscopeDSP.inl:
 * 310-385): scans a block and appends trigger positions (steady-clock sample
 * indices) per stream.  states / peaks / counts are HOST arrays; a stream that
 * produces more than max_peaks keeps the first max_peaks and reports the full
 * count. */
typedef struct sgz_scope_trigger_state { double state; uint64_t cross_origin; int32_t is_peak_holding; int32_t reserved; } sgz_scope_trigger_state;
int  sgz_scope_trigger(sgz_context* ctx, int mode, int channel, const float* audio, int mem, int layout,
                       uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride, double threshold,
                       double hysteresis, uint64_t steady_clock, sgz_scope_trigger_state* states,
                       uint64_t* peaks, uint32_t* counts, uint32_t max_peaks);
/* The same with states / peaks / counts (and the audio) in DEVICE memory: the
 * call only enqueues the kernel on the context's stream - no copy, no wait -
 * so a device-resident pipeline keeps the kernel's rate (the host-array form
 * above spends most of its time copying the peak lists back). */
int  sgz_scope_trigger_device(sgz_context* ctx, int mode, int channel, const float* d_audio, int layout,
                              uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride, double threshold,
                              double hysteresis, uint64_t steady_clock, sgz_scope_trigger_state* d_states,
                              uint64_t* d_peaks, uint32_t* d_counts, uint32_t max_peaks);

/* Replaces `TriggeringProcessor::update` and `::processMutating`
 * (Source/Oscilloscope/StreamPreprocessing.h:54-75, :78-206): the consumer of
 * the trigger positions above.  Per stream and callback block it decides how
 * many samples go into the back buffer before each window, and which window
 * (first sample, length) is swapped to the front buffer once the samples
 * around a trigger position are complete.  Integer bookkeeping on the host,
 * bit-exact; the reference's `audioProcessing(..., channelData.back)` (:91) and
 * `channelData.swapBuffers(cappedSize, -bufferedSamples)` (:194) are reported
 * as events for the caller to act on, in order:
 *   SGZ_WINDOW_PROCESS  `size` samples starting at steady clock `origin` are
 *                       appended to the back buffer (`buffered` = samples the
 *                       back buffer held before them);
 *   SGZ_WINDOW_SWAP     the `size` samples starting at clock `origin`
 *                       (= frontOrigin) become the displayed window; `buffered`
 *                       = bufferedSamples, i.e. swapBuffers' offset is -buffered.
 * states: HOST [stream] in/out (zero-initialise; set window_size /
 * window_changed as `setSettings` does, :45-52).  peaks: HOST
 * [stream][max_peaks], the queue front first; peak_counts: in = queued
 * positions (append what sgz_scope_trigger found), out = positions left, which
 * are moved to the front of the stream's row.  events: HOST
 * [stream][max_events]; event_counts: the number produced (a stream that
 * needs more than max_events returns SGZ_EINVAL with its state untouched).
 * steady_clock: the block's first sample (ctx.getPlayhead().getSteadyClock()). */
enum { SGZ_WINDOW_PROCESS = 0, SGZ_WINDOW_SWAP = 1 };
typedef struct sgz_scope_window_state {
    double   window_size;                 /* windowSize */
    uint64_t old_peak, current_peak, buffered_samples, front_origin, steady_clock;
    int32_t  is_working_on_peak, window_changed;
} sgz_scope_window_state;
typedef struct sgz_scope_window_event { uint64_t origin, size, buffered; uint32_t kind, reserved; } sgz_scope_window_event;
int  sgz_scope_trigger_windows(uint32_t n_streams, sgz_scope_window_state* states, uint64_t* peaks, uint32_t* peak_counts,
                               uint32_t max_peaks, uint64_t steady_clock, uint64_t n_samples,
                               sgz_scope_window_event* events, uint32_t* event_counts, uint32_t max_events);
/* Replaces `calculateFundamentalPeriod` + `calculateTriggeringOffset`
 * (Source/Oscilloscope/OscilloscopeDSP.inl:61-308, Spectral trigger mode):
 * 8192-point float64 FFT of the newest samples, fundamental pick with
 * harmonic rejection and quadratic bin offset, 8-deep median, then a Goertzel
 * phase at the fractional bin -> sample offset.  state / result: HOST [stream]. */
typedef struct sgz_scope_bin_record { uint64_t index; double value; double offset; } sgz_scope_bin_record;
typedef struct sgz_scope_period_state { sgz_scope_bin_record median[8]; uint64_t median_pos; } sgz_scope_period_state;
typedef struct sgz_scope_period_params { uint32_t abi_version; int32_t channel; double effective_window_size, sample_rate, threshold, hysteresis, phase_offset_degrees; } sgz_scope_period_params;
typedef struct sgz_scope_period_result { sgz_scope_bin_record record; double fundamental, cycle_samples, phase, sample_offset; } sgz_scope_period_result;
int  sgz_scope_period(sgz_context* ctx, const sgz_scope_period_params* params, const float* audio, int mem,
                      int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
                      sgz_scope_period_state* states, sgz_scope_period_result* results);

/* ---- oscilloscope colouring and envelopes -------------
 * sgz_scope_colour replaces the per-sample work of
 * `Oscilloscope::StreamState::audioProcessing`
 * (Source/Oscilloscope/OscilloscopeDSP.inl:426-710) for every stereo pair:
 * the RMS envelope of the selected channel mode (:505-586, stored only in
 * SGZ_ENVELOPE_RMS mode, :682-699), the 3-band crossover at 300 / 3000 Hz
 * (`channelData.tuneCrossOver`, ChannelData.h:163-166), the one-pole
 * smoothing of the squared bands (`filterStates`, :462-470) for left, right,
 * mid (= left + right bands) and side, and `accumulateColour` (:474-497)
 * into four RGBA8 colour streams per pair (colourData of the two channels,
 * auxColourData for mid / side, :606-651).
 * states: HOST [stream][2] (index 0 = left / mid, 1 = right / side; the
 * ChannelState of ChannelData.h:68-77), in/out.  rgba: [stream][4][n][4]
 * bytes, planes in the order left, right, mid, side, r g b a per pixel.
 * envelope_gain: HOST [stream] `envelopeGain` (0 unless RMS mode), may be
 * NULL.
 * sgz_scope_peak_envelope replaces `Oscilloscope::runPeakFilter` (:712-886):
 * the largest |sample| of the window (whole groups of 8 samples, as the AVX
 * loop) per channel mode folded into the two envelope states with
 * `coeff` (:742-743, computed by the caller), and `state.autoGain`.
 * envelopes: HOST [stream][2] floats in/out; auto_gain: HOST [stream]. */
enum { SGZ_ENVELOPE_NONE = 0, SGZ_ENVELOPE_RMS = 1, SGZ_ENVELOPE_PEAK = 2 };
typedef struct sgz_scope_colour_config {
    uint32_t abi_version;
    int32_t  channel_mode;        /* SGZ_OSC_* (all six) */
    int32_t  envelope_mode;       /* SGZ_ENVELOPE_* */
    double   sample_rate;
    double   colour_smoothing_ms; /* content->colourSmoothing.getTransformedValue() */
    double   envelope_window;     /* content->envelopeWindow.getNormalizedValue() */
    double   colouring_blend;     /* content->frequencyColouringBlend */
    float    colours[3][3];       /* low, mid, high colour as r g b in 0..1 */
    uint8_t  key_left[4], key_right[4]; /* ChannelState::defaultKey of the two channels, r g b a */
} sgz_scope_colour_config;
typedef struct sgz_scope_colour_state {
    float smooth_filters[3];      /* ChannelState::smoothFilters */
    float aux_smooth_filter[3];   /* ChannelState::auxSmoothFilter */
    float network[24];            /* ChannelState::network: 12 second-order sections x 2 */
    float envelope;               /* ChannelState::envelope */
} sgz_scope_colour_state;
int  sgz_scope_colour(sgz_context* ctx, const sgz_scope_colour_config* cfg, const float* audio, int mem,
                      int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
                      sgz_scope_colour_state* states, uint8_t* rgba, int rgba_mem, double* envelope_gain);
int  sgz_scope_peak_envelope(sgz_context* ctx, int channel_mode, const float* audio, int mem, int layout,
                             uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride, double coeff,
                             float* envelopes, double* auto_gain);

/* ---- allocation-free forward FFT ------------------------------------------
 * No counterpart in the reference.  For callers that cannot allocate device
 * memory at call time (real-time audio / render callbacks, hosts that forbid
 * cudaMalloc after start-up, processes whose device pool is already spent):
 * the unnormalised forward DFT of `batch` complex sequences of `n` points
 * (n a power of two >= 2), natural order, X[k] = sum x[j] exp(-2 pi i j k / n),
 * like sgz_spectrum_transform but without a handle.  d_in and d_out are
 * DEVICE buffers of batch * n interleaved (re, im) floats, distinct, owned by
 * the caller.  The call never invokes cudaMalloc/cudaFree or any allocating
 * routine: its working buffer is a fixed 8 MiB array inside the library
 * (resident on every GPU that loads the library, until the process exits), so
 * n * batch must not exceed SGZ_NOALLOC_MAX_POINTS and batch must not exceed
 * 65535 (SGZ_EINVAL otherwise); d_in and d_out must not overlap anywhere.
 * The first call on each device looks up that array's address (with that
 * device current); make it before starting any stream capture.  All calls
 * share the one working buffer: calls that overlap on the GPU (different
 * streams, or different host threads) race with each other; issue them on one
 * stream or order them yourself.  Work is enqueued on `cuda_stream`
 * (a cudaStream_t, NULL = default stream) and is asynchronous. */
#define SGZ_NOALLOC_MAX_POINTS (1u << 20)
int  sgz_fft_forward_noalloc(const float* d_in, float* d_out, uint32_t n, uint32_t batch,
                             void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* SGZ_H */
