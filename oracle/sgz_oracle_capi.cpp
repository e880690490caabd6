// sgz_oracle_capi.cpp — C entry points over the CPU oracle (sgz_oracle.hpp) for ctypes.
//
// TEST INFRASTRUCTURE ONLY (see the header of sgz_oracle.hpp): used by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
// The restatement is pinned bit for bit by the reference's own source (oracle/_ref, see below); the cpl primitives under
// both are decisions (no reference tests or golden vectors exist; cpl is absent).
//
// The handle mirrors the product's batched C ABI (include/sgz.h) so one test
// harness drives both: `n_streams` TransformPair objects sharing one
// TransformConstant, as in SpectrumDSP.cpp:83-102.
//
// Built twice: libsgz_oracle.so over the restatement (sgz_oracle.hpp), and - with -DSGZ_ORACLE_REF, by
// oracle/build_ref.sh - oracle/_ref/libsgz_oracle_ref.so over the REFERENCE's own TransformDSP.inl / TransformPair.h /
// TransformConstant.h (ref_adapter.hpp); tests/test_oracle_ref.py holds the two to the same bits.
#ifdef SGZ_ORACLE_REF
#include "ref_adapter.hpp"
#else
#include "sgz_oracle.hpp"
#endif
#include "../include/sgz.h"

#include <atomic>
#include <limits>
#include <memory>
#include <thread>

#ifdef SGZ_ORACLE_REF
using namespace sgzref;
#else
using namespace sgzo;
namespace
{
	// first / second scalar of a result or state entry
	template<typename S> inline S first(const UComplexFilter<S>& v) { return v.a; }
	template<typename S> inline S second(const UComplexFilter<S>& v) { return v.b; }
	template<typename S> inline void assign(UComplexFilter<S>& v, S a, S b) { v.a = a; v.b = b; }
	template<typename T> inline ComplexResonator<T>& resonatorOf(TransformPair<T>& p) { return p.cresonator; }
}
#endif

namespace
{
	struct IOracle
	{
		virtual ~IOracle() = default;
		virtual void info(sgz_spectrum_info* out) = 0;
		virtual void window(double* out) = 0;
		virtual void freqs(double* out) = 0;
		virtual void slopes(double* out) = 0;
		virtual void binMap(int32_t* mode, int32_t* first, int32_t* count, int32_t* fallback) = 0;
		virtual int prepare(const float* audio, int layout, uint64_t n, uint64_t stride, uint64_t split) = 0;
		virtual int preparePrelim(const float* hist, uint64_t nh, uint64_t hstride, uint64_t split, const float* prelim, uint64_t np, uint64_t pstride, int layout) = 0;
		virtual void transform() = 0;
		virtual void map() = 0;
		virtual void post() = 0;
		virtual void rawFFT(double* out) = 0;
		virtual void transformResult(double* out) = 0;
		virtual void results(int graph, float* out) = 0;
		virtual void states(int graph, float* out) = 0;
		virtual void setStates(int graph, const float* in) = 0;
		virtual void clearLineGraphs() = 0;
		virtual void clearAudio() = 0;
		virtual int frames(const float* audio, int layout, uint64_t n, uint64_t stride, uint32_t hop, uint32_t nFrames, float* results, int threads, uint32_t firstStream, uint32_t numStreams) = 0;
		virtual int resonate(const float* audio, int layout, uint64_t n, uint64_t stride) = 0;
		virtual void resonatorState(double* out) = 0;
		virtual int audioEntry(const float* hist, uint64_t nh, uint64_t hstride, const float* block, uint64_t nb, uint64_t bstride, int layout, float* frames, uint32_t maxFrames, uint32_t* numFrames) = 0;
	};

	// de-interleave helper: returns planar pointers for stream s
	struct StreamAudio
	{
		std::vector<float> storage;
		const float* ch[2];
	};

	static void fetch(const float* audio, int layout, uint64_t n, uint64_t stride, std::size_t s, StreamAudio& out)
	{
		const float* base = audio + s * stride;
		if (layout == SGZ_LAYOUT_PLANAR)
		{
			out.ch[0] = base; out.ch[1] = base + n;
		}
		else
		{
			out.storage.resize(2 * n);
			for (uint64_t i = 0; i < n; ++i) { out.storage[i] = base[2 * i]; out.storage[n + i] = base[2 * i + 1]; }
			out.ch[0] = out.storage.data(); out.ch[1] = out.storage.data() + n;
		}
	}

	template<typename T>
	struct Oracle : IOracle
	{
		TransformConstant<T> constant;
		std::vector<TransformPair<T>> pairs;
		sgz_spectrum_config cfg;

		Oracle(const sgz_spectrum_config& c, uint32_t nStreams) : pairs(nStreams), cfg(c)
		{
			// order of derivation: Spectrum.cpp:365-394 (scalars), :412-431 (configuration, sample
			// rate), :489 setStorage, :571-576 remapFrequencies, :578-581 slope map, :583-588
			// window kernel, :590-595 resonator
			constant.sampleBufferSize = c.sample_buffer_size;
			constant.algo = static_cast<TransformAlgorithm>(c.algo);
			constant.binPolation = static_cast<BinInterpolation>(c.bin_polation);
			constant.displayMode = static_cast<DisplayMode>(c.display_mode);
			constant.lowDBs = c.low_dbs; constant.clipDB = c.clip_db; constant.highDBs = c.high_dbs;
			constant.filter[0].pole = static_cast<T>(c.pole[0]);
			constant.filter[1].pole = static_cast<T>(c.pole[1]);
			constant.configuration = static_cast<SpectrumChannels>(c.configuration);
			constant.sampleRate = static_cast<T>(c.sample_rate);
			std::size_t ts;
			constant.setStorage(c.axis_points, c.window_size, ts);
			constant.remapFrequencies(c.view_left, c.view_right, static_cast<ViewScaling>(c.view_scaling), c.min_freq);
			constant.generateSlopeMap(c.slope_a, c.slope_b);
			constant.regenerateWindowKernel(static_cast<WindowTypes>(c.window_type), static_cast<WindowShape>(c.window_shape));
			constant.remapResonator(c.free_q != 0, dsp::windowCoefficientCount(static_cast<WindowTypes>(c.window_type)));
			for (auto& p : pairs)
				for (auto& g : p.lineGraphs) g.resize(c.axis_points);
		}

		void info(sgz_spectrum_info* out) override
		{
			std::memset(out, 0, sizeof(*out));
			out->transform_size = (uint32_t)constant.transformSize;
			out->window_size = (uint32_t)constant.windowSize;
			out->axis_points = (uint32_t)constant.axisPoints;
			out->n_streams = (uint32_t)pairs.size();
			out->state_channels = (uint32_t)constant.getStateConfigurationChannels();
			out->window_kernel_scale = (float)constant.windowKernelScale;
			out->inv_size = (float)static_cast<T>(constant.windowKernelScale / (constant.windowSize * 0.5));
			const auto topFrequency = constant.sampleRate / 2;
			out->freq_to_bin = (float)static_cast<T>((constant.transformSize >> 1) / topFrequency);
			out->resonator_vectors = (uint32_t)constant.resonator.numVectors;
			out->device = -1;
			std::vector<int32_t> mode(constant.axisPoints), a(constant.axisPoints), b(constant.axisPoints), c(constant.axisPoints);
			binMap(mode.data(), a.data(), b.data(), c.data());
			uint32_t bp = (uint32_t)constant.axisPoints;
			for (std::size_t x = 0; x < constant.axisPoints; ++x) if (mode[x] == 1) { bp = (uint32_t)x; break; }
			out->break_point = bp;
		}
		void window(double* out) override { for (std::size_t i = 0; i < constant.transformSize; ++i) out[i] = i < constant.windowSize ? (double)constant.windowKernel[i] : 0.0; }
		void freqs(double* out) override { for (std::size_t i = 0; i < constant.axisPoints; ++i) out[i] = constant.mappedFrequencies[i]; }
		void slopes(double* out) override { for (std::size_t i = 0; i < constant.axisPoints; ++i) out[i] = constant.slopeMap[i]; }

		// The integer skeleton of the loops at TransformDSP.inl:562-639 (mono), :875-984
		// (Separate/MidSide), :666-850 (Phase), :995-1096 (Complex): which display points
		// interpolate, which take max-of-bins, and over which bins.
		void binMap(int32_t* mode, int32_t* first, int32_t* count, int32_t* fallback) override
		{
			typedef std::ptrdiff_t ssz;
			const int N = (int)constant.transformSize;
			const int P = (int)constant.axisPoints;
			const std::size_t numBins = N >> 1;
			const auto topFrequency = constant.sampleRate / 2;
			const auto freqToBin = static_cast<T>(numBins / topFrequency);
			const auto& mf = constant.mappedFrequencies;
			const bool isComplex = constant.configuration == SpectrumChannels::Complex;
			const bool isPhase = constant.configuration == SpectrumChannels::Phase;
			const double fftBandwidth = isComplex ? 1.0 / (numBins * 2) : 1.0 / numBins;
			auto interp = [&](int x)
			{
				mode[x] = 0; count[x] = 0; fallback[x] = 0;
				if (constant.binPolation == BinInterpolation::None)
				{
					std::size_t hi = isComplex ? (std::size_t)N : (isPhase ? numBins - 1 : numBins);
					first[x] = (int32_t)Math::confineTo<std::size_t>((std::size_t)(mf[x] * freqToBin + 0.5), 0, hi);
				}
				else
					first[x] = (int32_t)std::floor(mf[x] * freqToBin);
			};
			auto maxOf = [&](int x, ssz& oldBin)
			{
				ssz bin = static_cast<ssz>(static_cast<std::size_t>(mf[x] * freqToBin));
				ssz diff = bin - oldBin;
				ssz counter = diff ? 1 : 0;
				mode[x] = 1; first[x] = (int32_t)(oldBin + counter);
				int32_t c = 0;
				do { c++; diff--; } while (diff > 0);
				count[x] = c;
				fallback[x] = isPhase ? 0 : (int32_t)bin;
				oldBin = bin;
			};
			int x = 0;
			ssz oldBin = 0;
			if (!isComplex)
			{
				for (x = 0; x < P - 1; ++x)
				{
					double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
					if (bwForLine > fftBandwidth) break;
					interp(x);
				}
				const bool neverBroke = x == P - 1;
				if (x < P) oldBin = static_cast<ssz>(mf[x] * freqToBin);
				for (; x < P; ++x) maxOf(x, oldBin);
				// Phase + Linear/Lanczos: bandWidthBreakingPoint stays at axisPoints when the loop
				// never breaks, so the magnitude pass (:709, :759) also interpolates the last point
				// and the max-of-bins loop does not run (:808-813)
				if (isPhase && neverBroke && constant.binPolation != BinInterpolation::None) interp(P - 1);
			}
			else
			{
				while (x < P)
				{
					for (; x < P; ++x)
					{
						if (x != P - 1)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth) break;
						}
						interp(x);
					}
					if (x != P) oldBin = static_cast<ssz>(mf[x] * freqToBin);
					for (; x < P; ++x)
					{
						if (x != P - 1)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine < fftBandwidth) break;
						}
						maxOf(x, oldBin);
					}
				}
			}
		}

		static AudioPair makeViews(const float* l, const float* r, uint64_t n, uint64_t split)
		{
			AudioPair v;
			if (split > n) split = n;
			// segment 0 = oldest `split` samples, segment 1 = the rest (a wrapped ring)
			v[0].seg[0] = l; v[0].len[0] = split; v[0].seg[1] = l + split; v[0].len[1] = n - split;
			v[1].seg[0] = r; v[1].len[0] = split; v[1].seg[1] = r + split; v[1].len[1] = n - split;
			return v;
		}

		int prepare(const float* audio, int layout, uint64_t n, uint64_t stride, uint64_t split) override
		{
			int rc = SGZ_OK;
			StreamAudio sa;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				fetch(audio, layout, n, stride, s, sa);
				if (!pairs[s].prepareTransform(constant, makeViews(sa.ch[0], sa.ch[1], n, split))) rc = SGZ_SKIPPED;
			}
			return rc;
		}
		int preparePrelim(const float* hist, uint64_t nh, uint64_t hstride, uint64_t split, const float* prelim, uint64_t np, uint64_t pstride, int layout) override
		{
			int rc = SGZ_OK;
			StreamAudio sh, sp;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				fetch(hist, layout, nh, hstride, s, sh);
				fetch(prelim, layout, np, pstride, s, sp);
				if (!pairs[s].prepareTransform(constant, makeViews(sh.ch[0], sh.ch[1], nh, split), { sp.ch[0], sp.ch[1] }, np)) rc = SGZ_SKIPPED;
			}
			return rc;
		}
		void transform() override { for (auto& p : pairs) p.doTransform(constant); }
		void map() override { for (auto& p : pairs) p.mapToLinearSpace(constant); }
		void post() override { for (auto& p : pairs) p.postProcessStdTransform(constant); }
		void rawFFT(double* out) override
		{
			const std::size_t N = constant.transformSize;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				auto* z = pairs[s].getRawFFT(constant);
				for (std::size_t i = 0; i < N; ++i) { out[2 * (s * N + i)] = z[i].real(); out[2 * (s * N + i) + 1] = z[i].imag(); }
			}
		}
		void transformResult(double* out) override
		{
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				std::size_t n;
				const T* r = pairs[s].getTransformResult(constant, &n);
				for (std::size_t i = 0; i < n; ++i) out[s * n + i] = r[i];
			}
		}
		void results(int graph, float* out) override
		{
			const std::size_t P = constant.axisPoints;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				const auto& r = pairs[s].lineGraphs[graph].getResults(P);
				for (std::size_t i = 0; i < P; ++i) { out[2 * (s * P + i)] = first(r[i]); out[2 * (s * P + i) + 1] = second(r[i]); }
			}
		}
		void states(int graph, float* out) override
		{
			const std::size_t P = constant.axisPoints;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				auto& g = pairs[s].lineGraphs[graph]; g.resize(P);
				for (std::size_t i = 0; i < P; ++i) { out[2 * (s * P + i)] = first(g.states[i]); out[2 * (s * P + i) + 1] = second(g.states[i]); }
			}
		}
		void setStates(int graph, const float* in) override
		{
			const std::size_t P = constant.axisPoints;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				auto& g = pairs[s].lineGraphs[graph]; g.resize(P);
				for (std::size_t i = 0; i < P; ++i) { assign(g.states[i], in[2 * (s * P + i)], in[2 * (s * P + i) + 1]); }
			}
		}
		void clearLineGraphs() override { for (auto& p : pairs) p.clearLineGraphStates(); }
		void clearAudio() override { for (auto& p : pairs) p.clearAudioState(); }

		// The body of the loop at SpectrumRendering.cpp:622-634 per frame and stream; the thread
		// pool over streams stands in for cpl::jobs::parallel_for (SpectrumDSP.cpp:83).
		int frames(const float* audio, int layout, uint64_t n, uint64_t stride, uint32_t hop, uint32_t nFrames, float* results, int threads, uint32_t firstStream, uint32_t numStreams) override
		{
			const bool rsnt = constant.algo == TransformAlgorithm::RSNT;
			const std::size_t W = constant.windowSize, P = constant.axisPoints;
			if (!rsnt && nFrames && n < (uint64_t)(nFrames - 1) * hop + W) return SGZ_EINVAL;
			if (rsnt && n < (uint64_t)nFrames * hop) return SGZ_EINVAL;
			if ((uint64_t)firstStream + numStreams > pairs.size()) return SGZ_EINVAL;
			std::atomic<std::size_t> next{ firstStream };
			const std::size_t end = (std::size_t)firstStream + numStreams;
			auto worker = [&]()
			{
				StreamAudio sa;
				for (;;)
				{
					const std::size_t s = next.fetch_add(1);
					if (s >= end) break;
					fetch(audio, layout, n, stride, s, sa);
					for (uint32_t f = 0; f < nFrames; ++f)
					{
						const uint64_t off = (uint64_t)f * hop;
						auto& p = pairs[s];
						if (rsnt)
						{
							// audioEntryPoint, RSNT branch (TransformDSP.inl:1178-1197): resonate one blob of
							// `sampleBufferSize` (= hop) samples, then addAudioFrame
							p.resonatingDispatch(constant, { sa.ch[0] + off, sa.ch[1] + off }, hop);
						}
						else
						{
							auto views = makeViews(sa.ch[0] + off, sa.ch[1] + off, W, W);
							if (!p.prepareTransform(constant, views)) continue;
							p.doTransform(constant);
						}
						p.mapToLinearSpace(constant);
						p.postProcessStdTransform(constant);
						if (results)
						{
							for (int g = 0; g < 2; ++g)
							{
								const auto& r = p.lineGraphs[g].getResults(P);
								float* o = results + (((s * nFrames + f) * 2 + g) * P) * 2;
								for (std::size_t i = 0; i < P; ++i) { o[2 * i] = first(r[i]); o[2 * i + 1] = second(r[i]); }
							}
						}
					}
				}
			};
			if (threads <= 1) worker();
			else
			{
				std::vector<std::thread> pool;
				for (int t = 0; t < threads; ++t) pool.emplace_back(worker);
				for (auto& t : pool) t.join();
			}
			return SGZ_OK;
		}

		int resonate(const float* audio, int layout, uint64_t n, uint64_t stride) override
		{
			StreamAudio sa;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				fetch(audio, layout, n, stride, s, sa);
				pairs[s].resonatingDispatch(constant, { sa.ch[0], sa.ch[1] }, n);
			}
			return SGZ_OK;
		}

		// raw resonator state [stream][channel][vector][filter] complex
		void resonatorState(double* out) override
		{
			const std::size_t ch = constant.getStateConfigurationChannels(), V = constant.resonator.numVectors, P = constant.axisPoints;
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				auto& r = resonatorOf(pairs[s]);
				r.ensure(constant.resonator);
				for (std::size_t k = 0; k < ch * V * P; ++k)
				{
					out[2 * (s * ch * V * P + k)] = r.real[k];
					out[2 * (s * ch * V * P + k) + 1] = r.imag[k];
				}
			}
		}

		// audioEntryPoint over one callback block (TransformDSP.inl:1165-1211); copies the
		// emitted sfbuf frames of every stream to frames[stream][frame][P][2].
		int audioEntry(const float* hist, uint64_t nh, uint64_t hstride, const float* block, uint64_t nb, uint64_t bstride, int layout, float* frames, uint32_t maxFrames, uint32_t* numFrames) override
		{
			StreamAudio sh, sb;
			const std::size_t P = constant.axisPoints;
			uint32_t emitted = 0;
			const auto authorityCounter = pairs.empty() ? 0 : pairs[0].processedSamplesSinceLastFrame; // SpectrumDSP.cpp:81
			for (std::size_t s = 0; s < pairs.size(); ++s)
			{
				fetch(block, layout, nb, bstride, s, sb);
				AudioPair views;
				if (constant.algo == TransformAlgorithm::FFT)
				{
					fetch(hist, layout, nh, hstride, s, sh);
					views = makeViews(sh.ch[0], sh.ch[1], nh, nh);
				}
				pairs[s].processedSamplesSinceLastFrame = authorityCounter; // SpectrumDSP.cpp:93
				pairs[s].audioEntryPoint(constant, &views, { sb.ch[0], sb.ch[1] }, nb);
				emitted = (uint32_t)pairs[s].sfbuf.size();
				for (uint32_t f = 0; f < emitted && f < maxFrames; ++f)
				{
					float* o = frames + ((s * maxFrames + f) * P) * 2;
					for (std::size_t i = 0; i < P; ++i) { o[2 * i] = first(pairs[s].sfbuf[f][i]); o[2 * i + 1] = second(pairs[s].sfbuf[f][i]); }
				}
				pairs[s].sfbuf.clear();
			}
			*numFrames = emitted;
			return SGZ_OK;
		}
	};
}

extern "C"
{
	void* sgzo_create(const sgz_spectrum_config* cfg, uint32_t n_streams, int use_double)
	{
		if (!cfg || cfg->abi_version != SGZ_ABI_VERSION || cfg->axis_points < 2 || cfg->window_size < 1) return nullptr;
		if (use_double) return static_cast<IOracle*>(new Oracle<double>(*cfg, n_streams));
		return static_cast<IOracle*>(new Oracle<float>(*cfg, n_streams));
	}
	void sgzo_destroy(void* h) { delete static_cast<IOracle*>(h); }
	void sgzo_get_info(void* h, sgz_spectrum_info* out) { static_cast<IOracle*>(h)->info(out); }
	void sgzo_get_window_kernel(void* h, double* out) { static_cast<IOracle*>(h)->window(out); }
	void sgzo_get_mapped_frequencies(void* h, double* out) { static_cast<IOracle*>(h)->freqs(out); }
	void sgzo_get_slope_map(void* h, double* out) { static_cast<IOracle*>(h)->slopes(out); }
	void sgzo_get_bin_map(void* h, int32_t* mode, int32_t* first, int32_t* count, int32_t* fallback) { static_cast<IOracle*>(h)->binMap(mode, first, count, fallback); }
	int sgzo_prepare(void* h, const float* audio, int layout, uint64_t n, uint64_t stride, uint64_t split) { return static_cast<IOracle*>(h)->prepare(audio, layout, n, stride, split); }
	int sgzo_prepare_preliminary(void* h, const float* hist, uint64_t nh, uint64_t hstride, uint64_t split, const float* prelim, uint64_t np, uint64_t pstride, int layout)
	{
		return static_cast<IOracle*>(h)->preparePrelim(hist, nh, hstride, split, prelim, np, pstride, layout);
	}
	void sgzo_transform(void* h) { static_cast<IOracle*>(h)->transform(); }
	void sgzo_map(void* h) { static_cast<IOracle*>(h)->map(); }
	void sgzo_postprocess(void* h) { static_cast<IOracle*>(h)->post(); }
	void sgzo_get_raw_fft(void* h, double* out) { static_cast<IOracle*>(h)->rawFFT(out); }
	void sgzo_get_transform_result(void* h, double* out) { static_cast<IOracle*>(h)->transformResult(out); }
	void sgzo_get_results(void* h, int graph, float* out) { static_cast<IOracle*>(h)->results(graph, out); }
	void sgzo_get_states(void* h, int graph, float* out) { static_cast<IOracle*>(h)->states(graph, out); }
	void sgzo_set_states(void* h, int graph, const float* in) { static_cast<IOracle*>(h)->setStates(graph, in); }
	void sgzo_clear_line_graph_states(void* h) { static_cast<IOracle*>(h)->clearLineGraphs(); }
	void sgzo_clear_audio_state(void* h) { static_cast<IOracle*>(h)->clearAudio(); }
	int sgzo_process_frames(void* h, const float* audio, int layout, uint64_t n, uint64_t stride, uint32_t hop, uint32_t n_frames, float* results, int threads, uint32_t first_stream, uint32_t num_streams)
	{
		return static_cast<IOracle*>(h)->frames(audio, layout, n, stride, hop, n_frames, results, threads, first_stream, num_streams);
	}
	void sgzo_get_resonator_state(void* h, double* out) { static_cast<IOracle*>(h)->resonatorState(out); }
	int sgzo_resonate(void* h, const float* audio, int layout, uint64_t n, uint64_t stride) { return static_cast<IOracle*>(h)->resonate(audio, layout, n, stride); }
	int sgzo_audio_entry_point(void* h, const float* hist, uint64_t nh, uint64_t hstride, const float* block, uint64_t nb, uint64_t bstride, int layout, float* frames, uint32_t max_frames, uint32_t* num_frames)
	{
		return static_cast<IOracle*>(h)->audioEntry(hist, nh, hstride, block, nb, bstride, layout, frames, max_frames, num_frames);
	}
	unsigned sgzo_hardware_concurrency(void) { return std::thread::hardware_concurrency(); }
}

#ifndef SGZ_ORACLE_REF // (the reference build covers the Spectrum path only)
// ---------------------------------------------------------------------------------------------
// secondary paths (sgz_oracle_scope.hpp)
// ---------------------------------------------------------------------------------------------
#include "sgz_oracle_scope.hpp"

namespace
{
	VScopeParams toParams(const sgz_vscope_config* c)
	{
		VScopeParams p;
		p.polar = c->polar != 0; p.scalePolar = c->scale_polar != 0; p.gain = c->gain; p.width = c->width; p.height = c->height;
		return p;
	}
	std::vector<float> channelOf(int channel, const StreamAudio& sa, uint64_t n)
	{
		std::vector<float> x(n);
		for (uint64_t i = 0; i < n; ++i) x[i] = oscChannel(static_cast<OscChannels>(channel), sa.ch[0][i], sa.ch[1][i]);
		return x;
	}
}

extern "C"
{
	void sgzo_vscope_vertices(const sgz_vscope_config* cfg, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, float* xyz)
	{
		const VScopeParams p = toParams(cfg);
		StreamAudio sa;
		for (uint32_t s = 0; s < S; ++s)
		{
			fetch(audio, layout, n, stride, s, sa);
			for (uint64_t i = 0; i < n; ++i)
			{
				const Vertex v = vscopeVertex(p, sa.ch[0][i], sa.ch[1][i], i, n);
				float* o = xyz + (uint64_t(s) * n + i) * 3;
				o[0] = v.x; o[1] = v.y; o[2] = v.z;
			}
		}
	}
	void sgzo_vscope_histogram(const sgz_vscope_config* cfg, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, uint32_t* hist)
	{
		const VScopeParams p = toParams(cfg);
		StreamAudio sa;
		std::fill(hist, hist + uint64_t(S) * p.width * p.height, 0u);
		for (uint32_t s = 0; s < S; ++s)
		{
			fetch(audio, layout, n, stride, s, sa);
			vscopeHistogram(p, sa.ch[0], sa.ch[1], n, hist + uint64_t(s) * p.width * p.height);
		}
	}
	void sgzo_vscope_meters(const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, float envelopeCoeff, float stereoCoeff, float speed, sgz_vscope_filters* filters)
	{
		StreamAudio sa;
		for (uint32_t s = 0; s < S; ++s)
		{
			fetch(audio, layout, n, stride, s, sa);
			VScopeFilters f;
			std::memcpy(static_cast<void*>(&f), &filters[s], sizeof(f));
			vscopeMeters(f, sa.ch[0], sa.ch[1], n, envelopeCoeff, stereoCoeff, speed);
			std::memcpy(&filters[s], static_cast<const void*>(&f), sizeof(f));
		}
	}
	uint64_t sgzo_scope_resample_count(const sgz_scope_resample_config* cfg)
	{
		ResampleParams p; p.left = cfg->left; p.right = cfg->right; p.inc = cfg->inc;
		return resampleCount(p);
	}
	void sgzo_scope_resample(const sgz_scope_resample_config* cfg, const double* samplePos, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, float* out)
	{
		StreamAudio sa;
		ResampleParams p; p.left = cfg->left; p.right = cfg->right; p.inc = cfg->inc; p.samplesPerPixel = cfg->samples_per_pixel;
		const uint64_t count = resampleCount(p);
		for (uint32_t s = 0; s < S; ++s)
		{
			fetch(audio, layout, n, stride, s, sa);
			const auto x = channelOf(cfg->channel, sa, n);
			p.samplePos = samplePos[s];
			lanczosResample(p, x.data(), (std::ptrdiff_t)n, out + uint64_t(s) * count * 2);
		}
	}
	void sgzo_scope_trigger(int mode, int channel, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, double threshold, double hysteresis,
		uint64_t steadyClock, sgz_scope_trigger_state* states, uint64_t* peaks, uint32_t* counts, uint32_t maxPeaks)
	{
		StreamAudio sa;
		for (uint32_t s = 0; s < S; ++s)
		{
			fetch(audio, layout, n, stride, s, sa);
			const auto x = channelOf(channel, sa, n);
			TriggerState st; st.state = states[s].state; st.isPeakHolding = states[s].is_peak_holding != 0; st.crossOrigin = states[s].cross_origin;
			std::vector<uint64_t> found;
			if (mode == 0) zeroCrossingTrigger(st, x.data(), n, threshold, steadyClock, found);
			else peakHoldTrigger(st, x.data(), n, threshold, hysteresis, steadyClock, found);
			states[s].state = st.state; states[s].is_peak_holding = st.isPeakHolding ? 1 : 0; states[s].cross_origin = st.crossOrigin;
			counts[s] = (uint32_t)found.size();
			for (uint32_t k = 0; k < found.size() && k < maxPeaks; ++k) peaks[uint64_t(s) * maxPeaks + k] = found[k];
		}
	}
	// TriggeringProcessor::update + processMutating per stream; same argument meaning as sgz_scope_trigger_windows
	int sgzo_scope_trigger_windows(uint32_t S, sgz_scope_window_state* states, uint64_t* peaks, uint32_t* peakCounts, uint32_t maxPeaks, uint64_t steadyClock,
		uint64_t n, sgz_scope_window_event* events, uint32_t* eventCounts, uint32_t maxEvents)
	{
		for (uint32_t s = 0; s < S; ++s)
		{
			TriggeringProcessor tp;
			const sgz_scope_window_state& in = states[s];
			tp.windowSize = in.window_size; tp.windowChanged = in.window_changed != 0;
			tp.oldPeak = in.old_peak; tp.currentPeak = in.current_peak; tp.bufferedSamples = in.buffered_samples;
			tp.frontOrigin = in.front_origin; tp.steadyClock = in.steady_clock; tp.isWorkingOnPeak = in.is_working_on_peak != 0;
			uint64_t* row = peaks + size_t(s) * maxPeaks;
			tp.peaks.assign(row, row + peakCounts[s]);
			std::vector<WindowEvent> found;
			tp.update(steadyClock);
			tp.processMutating(static_cast<std::size_t>(n), found);
			if (found.size() > maxEvents) return SGZ_EINVAL;
			for (size_t k = 0; k < found.size(); ++k)
				events[size_t(s) * maxEvents + k] = sgz_scope_window_event{ found[k].origin, found[k].size, found[k].buffered, uint32_t(found[k].kind), 0u };
			eventCounts[s] = uint32_t(found.size());
			sgz_scope_window_state out{};
			out.window_size = tp.windowSize; out.window_changed = tp.windowChanged ? 1 : 0;
			out.old_peak = tp.oldPeak; out.current_peak = tp.currentPeak; out.buffered_samples = tp.bufferedSamples;
			out.front_origin = tp.frontOrigin; out.steady_clock = tp.steadyClock; out.is_working_on_peak = tp.isWorkingOnPeak ? 1 : 0;
			states[s] = out;
			peakCounts[s] = uint32_t(tp.peaks.size());
			std::copy(tp.peaks.begin(), tp.peaks.end(), row);
		}
		return SGZ_OK;
	}
	void sgzo_scope_period(const sgz_scope_period_params* prm, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride,
		sgz_scope_period_state* states, sgz_scope_period_result* results)
	{
		StreamAudio sa;
		PeriodParams p; p.effectiveWindowSize = prm->effective_window_size; p.sampleRate = prm->sample_rate; p.threshold = prm->threshold;
		p.hysteresis = prm->hysteresis; p.phaseOffsetDegrees = prm->phase_offset_degrees;
		for (uint32_t s = 0; s < S; ++s)
		{
			fetch(audio, layout, n, stride, s, sa);
			const auto x = channelOf(prm->channel, sa, n);
			PeriodState st;
			for (int k = 0; k < 8; ++k) { st.median[k].index = states[s].median[k].index; st.median[k].value = states[s].median[k].value; st.median[k].offset = states[s].median[k].offset; }
			st.medianPos = states[s].median_pos;
			const PeriodResult r = spectralPeriod(st, p, x.data(), (std::ptrdiff_t)n);
			for (int k = 0; k < 8; ++k) { states[s].median[k].index = st.median[k].index; states[s].median[k].value = st.median[k].value; states[s].median[k].offset = st.median[k].offset; }
			states[s].median_pos = st.medianPos;
			results[s].record.index = r.record.index; results[s].record.value = r.record.value; results[s].record.offset = r.record.offset;
			results[s].fundamental = r.fundamental; results[s].cycle_samples = r.cycleSamples; results[s].phase = r.phase; results[s].sample_offset = r.sampleOffset;
		}
	}
}

// blendAndDispatchSpectrums — Source/Spectrum/SpectrumDSP.cpp:110-206, restated over flat arrays:
// results [S][F][2][P][2] (LineMain .magnitude = intensity), pairs of one view = consecutive streams.
extern "C" void sgzo_colour_columns(const float* results, uint32_t S, uint32_t F, uint32_t P, uint32_t pairs, const float* ratios, const float* colours, uint8_t* rgba)
{
	const uint32_t views = S / pairs;
	std::vector<std::array<float, 3>> colourBuffer(P);
	for (uint32_t v = 0; v < views; ++v)
		for (uint32_t s = 0; s < F; ++s)
		{
			for (auto& c : colourBuffer) c = { 0.f, 0.f, 0.f };
			for (uint32_t p = 0; p < pairs; ++p)
			{
				const float* sca = colours + p * 18; // generateSpectrogramColourRotation(p)
				const float* input = results + (((uint64_t(v) * pairs + p) * F + s) * 2 + 0) * P * 2;
				for (uint32_t i = 0; i < P; ++i)
				{
					const auto intensity = input[2 * i];
					if (intensity < 0) continue;
					std::array<float, 3> colour{ 0.f, 0.f, 0.f };
					if (intensity < 0.999f)
					{
						float accumulatedSum = 0;
						for (std::size_t c = 1; c < 6; ++c)
						{
							const auto nextScale = ratios[c];
							accumulatedSum += nextScale;
							if (accumulatedSum >= intensity)
							{
								const auto min = accumulatedSum - nextScale;
								const auto max = accumulatedSum;
								const auto mix = (intensity - min) / (max - min);
								const auto imix = 1 - mix;
								for (int k = 0; k < 3; ++k) colour[k] = sca[(c - 1) * 3 + k] * imix + sca[c * 3 + k] * mix;
								break;
							}
						}
					}
					else
					{
						for (int k = 0; k < 3; ++k) colour[k] = sca[15 + k];
					}
					for (std::size_t c = 0; c < 3; ++c) colourBuffer[i][c] += (1 - colourBuffer[i][c]) * colour[c];
				}
			}
			uint8_t* px = rgba + ((uint64_t(v) * F + s) * P) * 4;
			for (uint32_t p = 0; p < P; ++p)
			{
				px[4 * p + 0] = static_cast<uint8_t>(colourBuffer[p][0] * 255);
				px[4 * p + 1] = static_cast<uint8_t>(colourBuffer[p][1] * 255);
				px[4 * p + 2] = static_cast<uint8_t>(colourBuffer[p][2] * 255);
				px[4 * p + 3] = 255;
			}
		}
}

// calculateSpectrumColourRatios — Source/Spectrum/Spectrum.cpp:226-246
extern "C" void sgzo_colour_ratios(const double* specRatios, float* out)
{
	double acc = 0.0, vals[5];
	for (int i = 0; i < 5; ++i) { vals[i] = std::max(0.0001, specRatios[i]); acc += vals[i]; }
	acc += std::numeric_limits<float>::epsilon();
	out[0] = 0;
	for (int i = 0; i < 5; ++i) out[i + 1] = static_cast<float>(vals[i] / acc);
}

// ---- frequency tracker (oracle/sgz_oracle_tracker.hpp) --------------------------------------------
#include "sgz_oracle_tracker.hpp"
extern "C" void sgzo_track_peaks(const sgz_spectrum_config* cfg, int graph, uint32_t S, uint32_t transformSize, double windowScale, const float* mappedFreq,
	const float* slopeMap, const float* results /*[S][P][2] or NULL*/, const float* rawFFT /*[S][N+1][2] or NULL*/, const double* mouseFraction,
	double smoothingMs, double fps, sgz_tracker_state* states, sgz_tracker_result* out)
{
	sgz_oracle::TrackerView v{};
	v.graph = graph; v.fft = cfg->algo == SGZ_ALGO_FFT; v.configuration = cfg->configuration; v.binPolation = cfg->bin_polation;
	v.viewScale = cfg->view_scaling; v.viewLeft = cfg->view_left; v.viewRight = cfg->view_right; v.minLogFreq = cfg->min_freq;
	v.sampleRate = cfg->sample_rate; v.transformSize = transformSize; v.windowSize = cfg->window_size; v.axisPoints = cfg->axis_points;
	v.windowScale = windowScale; v.lowDbs = cfg->low_dbs; v.highDbs = cfg->high_dbs; v.mappedFrequencies = mappedFreq; v.slopeMap = slopeMap;
	for (uint32_t s = 0; s < S; ++s)
		sgz_oracle::trackPeak(v, results ? results + size_t(s) * cfg->axis_points * 2 : nullptr, rawFFT ? rawFFT + size_t(s) * (size_t(transformSize) + 1) * 2 : nullptr,
			mouseFraction[s], smoothingMs, fps, states ? states + s : nullptr, out + s);
}

// ---- oscilloscope colouring and envelopes (oracle/sgz_oracle_colour.hpp) ----------------------------
#include "sgz_oracle_colour.hpp"
extern "C" void sgzo_scope_colour(const sgz_scope_colour_config* cfg, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride,
	sgz_scope_colour_state* states, uint8_t* rgba, double* gain)
{
	std::vector<float> l(n), r(n);
	for (uint32_t s = 0; s < S; ++s)
	{
		const float* base = audio + size_t(s) * stride;
		for (uint64_t i = 0; i < n; ++i)
		{
			if (layout == SGZ_LAYOUT_INTERLEAVED) { l[i] = base[2 * i]; r[i] = base[2 * i + 1]; }
			else { l[i] = base[i]; r[i] = base[n + i]; }
		}
		const double g = sgz_oracle::audioProcessing(*cfg, l.data(), r.data(), n, states + 2 * size_t(s), rgba + size_t(s) * 16 * n);
		if (gain) gain[s] = g;
	}
}
extern "C" void sgzo_scope_peak_envelope(int channelMode, const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, double coeff, float* env, double* gain)
{
	std::vector<float> l(n), r(n);
	for (uint32_t s = 0; s < S; ++s)
	{
		const float* base = audio + size_t(s) * stride;
		for (uint64_t i = 0; i < n; ++i)
		{
			if (layout == SGZ_LAYOUT_INTERLEAVED) { l[i] = base[2 * i]; r[i] = base[2 * i + 1]; }
			else { l[i] = base[i]; r[i] = base[n + i]; }
		}
		gain[s] = sgz_oracle::runPeakFilter(channelMode, l.data(), r.data(), n, coeff, env + 2 * size_t(s));
	}
}
extern "C" void sgzo_vscope_peak_envelope(const float* audio, int layout, uint32_t S, uint64_t n, uint64_t stride, float envelopeCoeff, double deltaTime,
	sgz_vscope_filters* filters, double* gain)
{
	std::vector<float> l(n), r(n);
	for (uint32_t s = 0; s < S; ++s)
	{
		const float* base = audio + size_t(s) * stride;
		for (uint64_t i = 0; i < n; ++i)
		{
			if (layout == SGZ_LAYOUT_INTERLEAVED) { l[i] = base[2 * i]; r[i] = base[2 * i + 1]; }
			else { l[i] = base[i]; r[i] = base[n + i]; }
		}
		sgz_oracle::runVectorscopePeakFilter(l.data(), r.data(), n, envelopeCoeff, deltaTime, filters[s].envelope, gain + s);
	}
}
// the crossover alone, for the tests that pin it against scipy: bands [n][3]
extern "C" void sgzo_scope_crossover(double sampleRate, const float* x, uint64_t n, float* bands)
{
	const sgz_oracle::CrossoverCoeffs c = sgz_oracle::designCrossover(sampleRate);
	float z[24] = {};
	for (uint64_t i = 0; i < n; ++i)
	{
		float b[3];
		sgz_oracle::runNetwork(c, z, x[i], b);
		bands[3 * i] = b[0]; bands[3 * i + 1] = b[1]; bands[3 * i + 2] = b[2];
	}
}
#endif // SGZ_ORACLE_REF
