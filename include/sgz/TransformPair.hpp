// sgz/TransformPair.hpp — C++17 host-side mirror of the reference's Spectrum interface over the C ABI.
//
// `sgz::TransformConstant` and `sgz::TransformPairBatch` keep the member names, argument meaning,
// call order and error behaviour of `Signalizer::TransformConstant<float>` /
// `Signalizer::TransformPair<float>` (reference: Source/Spectrum/TransformConstant.h:47-241,
// Source/Spectrum/TransformPair.h:50-246), batched over the stereo pairs of one view: where the
// reference loops `parallel_for(numChannels / 2, ...)` over pair objects (SpectrumDSP.cpp:83-102),
// one TransformPairBatch call covers every pair.  Header-only; link with -lsgz_cuda.
#pragma once

#include <complex>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../sgz.h"

namespace sgz
{
	// CommonSignalizer.h:495-539
	enum class SpectrumChannels : int { Left = SGZ_CH_LEFT, Right, Merge, Mid = Merge, Side, OffsetForMono = Side, Phase, Separate, MidSide, Complex, End };
	// SpectrumParameters.h:48-75
	enum LineGraphs { LineMain = SGZ_LINE_MAIN, LineSecond, LineEnd };
	enum class BinInterpolation : int { None = SGZ_BIN_NONE, Linear, Lanczos };
	enum class DisplayMode : int { LineGraph = SGZ_DISPLAY_LINEGRAPH, ColourSpectrum };
	enum class TransformAlgorithm : int { FFT = SGZ_ALGO_FFT, RSNT };
	enum class ViewScaling : int { Linear = SGZ_VIEW_LINEAR, Logarithmic };

	// CommonSignalizer.h:541-591 (same 8-byte layout)
	template<typename Scalar>
	union UComplexFilter
	{
		struct { Scalar real, imag; };
		struct { Scalar magnitude, phase; };
		struct { Scalar leftMagnitude, rightMagnitude; };
	};

	struct Error : std::runtime_error
	{
		int code;
		Error(int c, const char* msg) : std::runtime_error(std::string("sgz error ") + std::to_string(c) + ": " + msg), code(c) {}
	};
	inline void check(int rc) { if (rc < 0) throw Error(rc, sgz_last_error()); }

	/// The read-only per-view constants.  Public fields and setters mirror the reference's; the
	/// derived tables are built by the library when a TransformPairBatch is created from it.
	class TransformConstant
	{
	public:
		TransformConstant() { sgz_spectrum_default_config(&cfg); }

		// TransformConstant.h:81-92.  transformSize = max(32, nextPow2(effectiveWindowSize))
		void setStorage(std::size_t elements, std::size_t effectiveWindowSize, std::size_t& outputTransformSize)
		{
			cfg.axis_points = static_cast<uint32_t>(elements);
			cfg.window_size = static_cast<uint32_t>(effectiveWindowSize);
			std::size_t n = 1; while (n < effectiveWindowSize) n <<= 1;
			outputTransformSize = n < 32 ? 32 : n;
		}
		// TransformConstant.h:125-180
		void remapFrequencies(double viewLeft, double viewRight, ViewScaling scaling, double minFreq)
		{
			cfg.view_left = viewLeft; cfg.view_right = viewRight; cfg.view_scaling = static_cast<int>(scaling); cfg.min_freq = minFreq;
		}
		// TransformConstant.h:109-118: slope = b * f^a
		void generateSlopeMap(double a, double b) { cfg.slope_a = a; cfg.slope_b = b; }
		// TransformConstant.h:104-107
		void regenerateWindowKernel(int windowType, int windowShape) { cfg.window_type = windowType; cfg.window_shape = windowShape; }
		// TransformConstant.h:120-123
		void remapResonator(bool shouldHaveFreeQ, std::size_t /*numVectors: derived from the window*/) { cfg.free_q = shouldHaveFreeQ; }
		// TransformConstant.h:183-186
		std::size_t getStateConfigurationChannels() const noexcept { return cfg.configuration > SGZ_CH_SIDE ? 2 : 1; }

		void setAlgorithm(TransformAlgorithm a) { cfg.algo = static_cast<int>(a); }
		void setConfiguration(SpectrumChannels c) { cfg.configuration = static_cast<int>(c); }
		void setBinPolation(BinInterpolation b) { cfg.bin_polation = static_cast<int>(b); }
		void setDisplayMode(DisplayMode m) { cfg.display_mode = static_cast<int>(m); }
		void setSampleRate(float fs) { cfg.sample_rate = fs; }
		void setDynamicRange(double lowDBs, double highDBs, double clipDB) { cfg.low_dbs = lowDBs; cfg.high_dbs = highDBs; cfg.clip_db = clipDB; }
		void setPeakFilterPoles(float lineMain, float lineSecond) { cfg.pole[0] = lineMain; cfg.pole[1] = lineSecond; }
		void setSampleBufferSize(std::size_t blob) { cfg.sample_buffer_size = static_cast<uint32_t>(blob); }

		sgz_spectrum_config cfg;
	};

	/// `nPairs` TransformPair objects on one GPU.  Member names as TransformPair.h:99-184.
	class TransformPairBatch
	{
	public:
		typedef UComplexFilter<float> UComplex;

		TransformPairBatch(const TransformConstant& constant, std::size_t nPairs, int device = 0)
		{
			check(sgz_spectrum_create(&constant.cfg, static_cast<uint32_t>(nPairs), device, &handle));
			check(sgz_spectrum_get_info(handle, &info));
		}
		~TransformPairBatch() { sgz_spectrum_destroy(handle); }
		TransformPairBatch(const TransformPairBatch&) = delete;
		TransformPairBatch& operator=(const TransformPairBatch&) = delete;

		/// bool prepareTransform(const Constant&, const AudioPair&) — TransformDSP.inl:39-231.
		/// `audio`: per pair `numSamples` samples per channel, oldest first; false = skip the frame.
		bool prepareTransform(const float* audio, std::size_t numSamples, std::size_t pairStride, int layout = SGZ_LAYOUT_PLANAR, int mem = SGZ_MEM_HOST)
		{
			const int rc = sgz_spectrum_prepare(handle, audio, mem, layout, numSamples, pairStride);
			check(rc);
			return rc == SGZ_OK;
		}
		/// The preliminary-audio overload — TransformDSP.inl:234-484.
		bool prepareTransform(const float* history, std::size_t historySamples, std::size_t historyStride,
			const float* preliminaryAudio, std::size_t numSamples, std::size_t preliminaryStride, int layout = SGZ_LAYOUT_PLANAR, int mem = SGZ_MEM_HOST)
		{
			const int rc = sgz_spectrum_prepare_preliminary(handle, history, historySamples, historyStride, preliminaryAudio, numSamples, preliminaryStride, mem, layout);
			check(rc);
			return rc == SGZ_OK;
		}
		void doTransform() { check(sgz_spectrum_transform(handle)); }                 // TransformDSP.inl:487-502
		void mapToLinearSpace() { check(sgz_spectrum_map(handle)); }                  // TransformDSP.inl:506-1135
		void postProcessStdTransform() { check(sgz_spectrum_postprocess(handle)); }   // TransformDSP.inl:1438-1441

		/// getTransformResult — [pair][axisPoints * channels * 2]
		std::vector<float> getTransformResult()
		{
			std::vector<float> out(std::size_t(info.n_streams) * info.axis_points * info.state_channels * 2);
			check(sgz_spectrum_get_transform_result(handle, out.data(), SGZ_MEM_HOST));
			return out;
		}
		/// getRawFFT — [pair][transformSize]
		std::vector<std::complex<float>> getRawFFT()
		{
			std::vector<std::complex<float>> out(std::size_t(info.n_streams) * info.transform_size);
			check(sgz_spectrum_get_raw_fft(handle, reinterpret_cast<float*>(out.data()), SGZ_MEM_HOST));
			return out;
		}
		/// lineGraphs[graph].getResults(axisPoints) — [pair][axisPoints]
		std::vector<UComplex> getResults(LineGraphs graph)
		{
			std::vector<UComplex> out(std::size_t(info.n_streams) * info.axis_points);
			check(sgz_spectrum_get_results(handle, graph, reinterpret_cast<float*>(out.data()), SGZ_MEM_HOST));
			return out;
		}
		void clearLineGraphStates() { check(sgz_spectrum_clear_line_graph_states(handle)); }
		void clearAudioState() { check(sgz_spectrum_clear_audio_state(handle)); }

		/// Every pair's `numFrames` frames at `hop` (the addAudioFrame loop, TransformDSP.inl:1139-1148,
		/// 1172-1204): results [pair][frame][graph][axisPoints].
		void processFrames(const float* audio, std::size_t numSamples, std::size_t pairStride, std::size_t hop, std::size_t numFrames,
			UComplex* results, int layout = SGZ_LAYOUT_PLANAR, int audioMem = SGZ_MEM_HOST, int resultsMem = SGZ_MEM_HOST)
		{
			check(sgz_spectrum_process_frames(handle, audio, audioMem, layout, numSamples, pairStride, static_cast<uint32_t>(hop),
				static_cast<uint32_t>(numFrames), reinterpret_cast<float*>(results), resultsMem));
		}

		/// The same with a choice of what comes back: SGZ_OUT_LINE_MAIN = graph LineMain of every frame, what addAudioFrame keeps
		/// in sfbuf ([pair][frame][axisPoints]); SGZ_OUT_LAST_FRAME = both graphs of the newest frame ([pair][graph][axisPoints]).
		void processFrames(const float* audio, std::size_t numSamples, std::size_t pairStride, std::size_t hop, std::size_t numFrames,
			UComplex* results, int output, int layout, int audioMem = SGZ_MEM_HOST, int resultsMem = SGZ_MEM_HOST)
		{
			check(sgz_spectrum_process_frames_select(handle, audio, audioMem, layout, numSamples, pairStride, static_cast<uint32_t>(hop),
				static_cast<uint32_t>(numFrames), reinterpret_cast<float*>(results), resultsMem, output));
		}
		/// Sizes every lazily grown buffer once (a caller with a deadline reserves before its first callback).
		void reserve(std::size_t numSamples, std::size_t pairStride, std::size_t numFrames)
		{
			check(sgz_spectrum_reserve(handle, numSamples, pairStride, static_cast<uint32_t>(numFrames)));
		}
		/// resonatingDispatch - TransformDSP.inl:1215-1295 (tensor cores for blocks of 2048 samples and more)
		void resonatingDispatch(const float* audio, std::size_t numSamples, std::size_t pairStride, int layout = SGZ_LAYOUT_PLANAR, int mem = SGZ_MEM_HOST)
		{
			check(sgz_resonator_process(handle, audio, mem, layout, numSamples, pairStride));
		}

		/// Device-resident audio history (the cpl::AudioStream side of TransformDSP.inl:49,65-88): reserve once,
		/// append each callback block, then run the frames that fell due (audioEntryPoint's scheduling, :1165-1211).
		void historyReserve(std::size_t capacity) { check(sgz_spectrum_history_reserve(handle, capacity)); }
		void historyAppend(const float* block, std::size_t numSamples, std::size_t pairStride, int layout = SGZ_LAYOUT_PLANAR, int mem = SGZ_MEM_HOST)
		{
			check(sgz_spectrum_history_append(handle, block, mem, layout, numSamples, pairStride));
		}
		/// returns the number of frames written to `results` ([pair][frames][graph][axisPoints])
		std::size_t historyProcess(std::size_t hop, std::size_t maxFrames, UComplex* results, int resultsMem = SGZ_MEM_HOST)
		{
			uint32_t n = 0;
			check(sgz_spectrum_history_process(handle, static_cast<uint32_t>(hop), static_cast<uint32_t>(maxFrames), reinterpret_cast<float*>(results), resultsMem, &n));
			return n;
		}

		sgz_spectrum* handle{ nullptr };
		sgz_spectrum_info info{};
	};
}
