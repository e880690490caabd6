// sgz_tables.cpp — see sgz_tables.hpp.
#include "sgz_tables.hpp"

#include <algorithm>
#include <cmath>

namespace sgz
{
	static const double kPi = 3.1415926535897932384626433832795;

	std::size_t nextPow2Inc(std::size_t x)
	{
		std::size_t p = 1;
		while (p < x) p <<= 1;
		return p;
	}

	static double sincPi(double x)
	{
		if (x == 0) return 1.0;
		const double px = kPi * x;
		return std::sin(px) / px;
	}

	double lanczosKernel(double d, int a)
	{
		if (d <= -a || d >= a) return 0.0;
		return sincPi(d) * sincPi(d / a);
	}

	// cosine-sum window family (cpl is absent; family and scaling are documented decisions,
	// DESIGN.md "cpl decisions")
	static std::vector<double> cosineTerms(int type)
	{
		switch (type)
		{
		case SGZ_WIN_HANN: return { 0.5, 0.5 };
		case SGZ_WIN_HAMMING: return { 0.54, 0.46 };
		case SGZ_WIN_BLACKMAN: return { 0.42, 0.5, 0.08 };
		case SGZ_WIN_BLACKMAN_HARRIS: return { 0.35875, 0.48829, 0.14128, 0.01168 };
		case SGZ_WIN_NUTTALL: return { 0.355768, 0.487396, 0.144232, 0.012604 };
		case SGZ_WIN_FLATTOP: return { 0.21557895, 0.41663158, 0.277263158, 0.083578947, 0.006947368 };
		default: return { 1.0 };
		}
	}

	static bool isMono(int conf) { return conf <= SGZ_CH_SIDE; }

	int buildTables(const sgz_spectrum_config& cfg, Tables& t)
	{
		if (cfg.abi_version != SGZ_ABI_VERSION) return SGZ_EINVAL;
		if (cfg.axis_points < 2 || cfg.window_size < 1 || cfg.window_size > (1u << 24)) return SGZ_EINVAL;
		if (cfg.configuration < SGZ_CH_LEFT || cfg.configuration > SGZ_CH_COMPLEX) return SGZ_EINVAL;
		if (cfg.bin_polation < SGZ_BIN_NONE || cfg.bin_polation > SGZ_BIN_LANCZOS) return SGZ_EINVAL;
		if (cfg.algo != SGZ_ALGO_FFT && cfg.algo != SGZ_ALGO_RSNT) return SGZ_EINVAL;
		if (cfg.view_scaling != SGZ_VIEW_LINEAR && cfg.view_scaling != SGZ_VIEW_LOGARITHMIC) return SGZ_EINVAL;
		if (cfg.window_type < 0 || cfg.window_type >= SGZ_WIN_END) return SGZ_EINVAL;
		if (!(cfg.sample_rate > 0)) return SGZ_EINVAL;

		// setStorage, TransformConstant.h:81-92
		t.W = cfg.window_size;
		t.N = static_cast<uint32_t>(std::max<std::size_t>(32, nextPow2Inc(t.W)));
		t.P = cfg.axis_points;
		t.numBins = t.N >> 1;
		t.stateChannels = cfg.configuration > SGZ_CH_SIDE ? 2 : 1; // getStateConfigurationChannels, :183-186
		t.sampleRate = cfg.sample_rate;
		const uint32_t N = t.N, P = t.P, W = t.W, numBins = t.numBins;

		// remapFrequencies, TransformConstant.h:125-180
		t.mappedFreq.resize(P);
		{
			const double viewSize = cfg.view_right - cfg.view_left;
			const bool cplx = cfg.configuration == SGZ_CH_COMPLEX;
			if (cfg.view_scaling == SGZ_VIEW_LINEAR)
			{
				const double halfSampleRate = t.sampleRate * 0.5;
				const double complexFactor = cplx ? 2.0 : 1.0;
				const double freqPerPixel = halfSampleRate / (P - 1);
				for (uint32_t i = 0; i < P; ++i)
					t.mappedFreq[i] = static_cast<float>(complexFactor * cfg.view_left * halfSampleRate + complexFactor * viewSize * i * freqPerPixel);
			}
			else
			{
				const double sampleSize = double(P - 1);
				const double end = t.sampleRate / 2;
				for (uint32_t i = 0; i < P; ++i)
				{
					if (!cplx)
						t.mappedFreq[i] = static_cast<float>(cfg.min_freq * std::pow(end / cfg.min_freq, cfg.view_left + viewSize * (i / sampleSize)));
					else
					{
						double arg = cfg.view_left + viewSize * i / sampleSize;
						if (arg < 0.5)
							t.mappedFreq[i] = static_cast<float>(cfg.min_freq * std::pow(end / cfg.min_freq, arg * 2));
						else
						{
							arg -= 0.5;
							const double power = cfg.min_freq * std::pow(end / cfg.min_freq, 1.0 - arg * 2);
							t.mappedFreq[i] = static_cast<float>(end + (end - power));
						}
					}
				}
			}
		}

		// generateSlopeMap, TransformConstant.h:109-118
		t.slope.resize(P);
		{
			const float a = static_cast<float>(cfg.slope_a), b = static_cast<float>(cfg.slope_b);
			for (uint32_t i = 0; i < P; ++i)
				t.slope[i] = b * static_cast<float>(std::pow(t.mappedFreq[i], a));
		}

		// regenerateWindowKernel, TransformConstant.h:104-107
		t.window.assign(N, 0.0f);
		const auto terms = cosineTerms(cfg.window_type);
		{
			const double D = cfg.window_shape == SGZ_WINSHAPE_PERIODIC ? double(W) : double(W > 1 ? W - 1 : 1);
			const double twoPi = 6.283185307179586476925286766559;
			double sum = 0;
			for (uint32_t n = 0; n < W; ++n)
			{
				double w = 0, sign = 1;
				for (std::size_t k = 0; k < terms.size(); ++k)
				{
					w += sign * terms[k] * std::cos(twoPi * double(k) * double(n) / D);
					sign = -sign;
				}
				t.window[n] = static_cast<float>(w);
				sum += static_cast<double>(t.window[n]);
			}
			t.windowKernelScale = static_cast<float>(sum > 0 ? double(W) / sum : 1.0);
			// the same window as a polynomial in cos(2 pi n / N) (Chebyshev T_k -> powers), for kernels that would rather
			// evaluate it than read it: only when the window spans the transform in its periodic shape
			t.winPolyOk = W == N && cfg.window_shape == SGZ_WINSHAPE_PERIODIC && terms.size() <= 5;
			if (t.winPolyOk)
			{
				static const double cheb[5][5] = { { 1, 0, 0, 0, 0 }, { 0, 1, 0, 0, 0 }, { -1, 0, 2, 0, 0 }, { 0, -3, 0, 4, 0 }, { 1, 0, -8, 0, 8 } };
				double p[5] = { 0, 0, 0, 0, 0 }, sg = 1;
				for (std::size_t k = 0; k < terms.size(); ++k, sg = -sg)
					for (int j = 0; j < 5; ++j) p[j] += sg * terms[k] * cheb[k][j];
				for (int j = 0; j < 5; ++j) t.winPoly[j] = static_cast<float>(p[j]);
			}
		}

		// scalars of mapToLinearSpace, TransformDSP.inl:522-540
		t.topFrequency = t.sampleRate / 2;
		t.freqToBin = static_cast<float>(numBins / t.topFrequency);
		t.invSize = static_cast<float>(t.windowKernelScale / (W * 0.5));

		// scalars of mapAndTransformDFTFilters, TransformDSP.inl:1308-1314, :1398
		{
			const double lowerFraction = std::pow(10.0, cfg.low_dbs / 20.0);
			const double upperFraction = std::pow(10.0, cfg.high_dbs / 20.0);
			t.deltaYRecip = static_cast<float>(1.0 / std::log(upperFraction / lowerFraction));
			t.minFracRecip = static_cast<float>(1.0 / lowerFraction);
			t.lowerClip = static_cast<float>(cfg.clip_db);
			for (int k = 0; k < 2; ++k)
			{
				t.pole[k] = cfg.pole[k];
				t.phasePole[k] = static_cast<float>(std::pow(cfg.pole[k], 0.3f));
			}
		}

		// ---- bin map: integer skeleton of the mapping loops ----
		t.mode.assign(P, 0); t.first.assign(P, 0); t.count.assign(P, 0); t.fallback.assign(P, 0);
		t.posL.assign(P, 0); t.posR.assign(P, 0); t.baseL.assign(P, 0); t.baseR.assign(P, 0);
		t.wL.assign(std::size_t(P) * kLanczosStride, 0); t.wR.assign(std::size_t(P) * kLanczosStride, 0);
		t.phaseNormPos.assign(P, 0); t.phaseNormFinal = 0;
		const bool cplx = cfg.configuration == SGZ_CH_COMPLEX;
		const bool phase = cfg.configuration == SGZ_CH_PHASE;
		const double fftBandwidth = cplx ? 1.0 / (numBins * 2) : 1.0 / numBins;
		const auto& mf = t.mappedFreq;
		auto bandwidthOf = [&](int x) -> double { return (mf[x + 1] - mf[x]) / t.topFrequency; };
		auto markInterp = [&](int x)
		{
			t.mode[x] = 0;
			const float pos = mf[x] * t.freqToBin;
			if (cfg.bin_polation == SGZ_BIN_NONE)
			{
				const std::size_t hi = cplx ? std::size_t(N) : (phase ? std::size_t(numBins - 1) : std::size_t(numBins));
				std::size_t idx = (std::size_t)(pos + 0.5);
				t.first[x] = static_cast<int32_t>(std::min(idx, hi));
			}
			else
				t.first[x] = static_cast<int32_t>(std::floor(pos));
		};
		std::ptrdiff_t oldBin = 0;
		auto markMax = [&](int x)
		{
			const std::ptrdiff_t bin = static_cast<std::ptrdiff_t>(static_cast<std::size_t>(mf[x] * t.freqToBin));
			std::ptrdiff_t diff = bin - oldBin;
			const std::ptrdiff_t counter = diff ? 1 : 0;
			t.mode[x] = 1;
			t.first[x] = static_cast<int32_t>(oldBin + counter);
			t.count[x] = static_cast<int32_t>(diff > 0 ? diff : 1);
			t.fallback[x] = phase ? 0 : static_cast<int32_t>(bin);
			oldBin = bin;
		};
		{
			int x = 0;
			const int iP = static_cast<int>(P);
			if (!cplx)
			{
				for (x = 0; x < iP - 1; ++x)
				{
					if (bandwidthOf(x) > fftBandwidth) break;
					markInterp(x);
				}
				t.breakPoint = static_cast<uint32_t>(x);
				if (x < iP) oldBin = static_cast<std::ptrdiff_t>(mf[x] * t.freqToBin);
				for (; x < iP; ++x) markMax(x);
			}
			else
			{
				t.breakPoint = P;
				while (x < iP)
				{
					for (; x < iP; ++x)
					{
						if (x != iP - 1 && bandwidthOf(x) > fftBandwidth) break;
						markInterp(x);
					}
					if (x != iP)
					{
						oldBin = static_cast<std::ptrdiff_t>(mf[x] * t.freqToBin);
						if (t.breakPoint == P) t.breakPoint = static_cast<uint32_t>(x);
					}
					for (; x < iP; ++x)
					{
						if (x != iP - 1 && bandwidthOf(x) < fftBandwidth) break;
						markMax(x);
					}
				}
			}
		}

		// interpolation positions and Lanczos weights (float32 position arithmetic as in
		// `lanczosFilter(csf, mf[x] * freqToBin, 5)` and `N - (mf[x] * freqToBin)`, :911-912)
		for (uint32_t x = 0; x < P; ++x)
		{
			const float pl = mf[x] * t.freqToBin;
			const float pr = static_cast<int>(N) - pl;
			t.posL[x] = pl; t.posR[x] = pr;
			const std::ptrdiff_t bl = static_cast<std::ptrdiff_t>(std::floor(pl)) - kLanczosSize + 1;
			const std::ptrdiff_t br = static_cast<std::ptrdiff_t>(std::floor(pr)) - kLanczosSize + 1;
			t.baseL[x] = static_cast<int32_t>(bl); t.baseR[x] = static_cast<int32_t>(br);
			for (int k = 0; k < kLanczosTaps; ++k)
			{
				t.wL[std::size_t(x) * kLanczosStride + k] = static_cast<float>(lanczosKernel(double(pl) - double(bl + k), kLanczosSize));
				t.wR[std::size_t(x) * kLanczosStride + k] = static_cast<float>(lanczosKernel(double(pr) - double(br + k), kLanczosSize));
			}
		}

		// Phase: progress of the lazy magnitude normalisation (TransformDSP.inl:700-727, :751-776)
		if (phase && cfg.bin_polation != SGZ_BIN_NONE)
		{
			// the loop `for (x = 0; x < axisPoints - 1; ++x)` leaves bandWidthBreakingPoint at
			// axisPoints when it never breaks; breakPoint == P-1 is exactly that case
			const std::size_t bwbp = t.breakPoint < P - 1 ? t.breakPoint : P;
			const std::size_t filterSize = cfg.bin_polation == SGZ_BIN_LANCZOS ? std::size_t(kLanczosSize) : 1;
			std::size_t normalizedPosition = 0;
			for (std::size_t x = 0; x < bwbp && x < P; ++x)
			{
				const float binPosition = mf[x] * t.freqToBin;
				while ((binPosition + static_cast<float>(filterSize)) > static_cast<float>(normalizedPosition) && x < bwbp - (filterSize + 1))
					normalizedPosition++;
				t.phaseNormPos[x] = static_cast<int32_t>(std::min<std::size_t>(normalizedPosition, N + 1));
			}
			t.phaseNormFinal = static_cast<int32_t>(std::min<std::size_t>(normalizedPosition, N + 1));
			if (bwbp == P) { markInterp(int(P) - 1); t.count[P - 1] = 0; t.fallback[P - 1] = 0; } // the magnitude pass covers the last point too
		}

		// ---- fused-kernel eligibility and the magnitude/squared split ----
		{
			const uint32_t M = numBins;
			uint32_t L = 1;
			const bool cplx = cfg.configuration == SGZ_CH_COMPLEX; // one channel over all N bins, magnitudes throughout
			bool ok = cfg.algo == SGZ_ALGO_FFT && (isMono(cfg.configuration) || cplx || cfg.configuration == SGZ_CH_SEPARATE || cfg.configuration == SGZ_CH_MIDSIDE
				|| cfg.configuration == SGZ_CH_PHASE);
			ok = ok && (N == 1024 || N == 2048 || N == 4096 || N == 8192 || N == 16384);
			const bool mono = isMono(cfg.configuration);
			const bool single = mono || cplx;
			auto tapL = [&](std::ptrdiff_t i)
			{
				if (i < 0 || i > std::ptrdiff_t(N)) return;
				if (mono && i >= std::ptrdiff_t(M) && cfg.bin_polation != SGZ_BIN_NONE) ok = false; // complex / raw entries
				L = std::max<uint32_t>(L, i <= std::ptrdiff_t(M) ? uint32_t(i) + 1 : M + 1);
			};
			auto tapR = [&](std::ptrdiff_t i)
			{
				if (i < 0 || i > std::ptrdiff_t(N)) return;
				L = std::max<uint32_t>(L, i >= std::ptrdiff_t(M) ? uint32_t(N - i) + 1 : M + 1);
			};
			for (uint32_t x = 0; x < P; ++x)
			{
				if (t.mode[x] == 0)
				{
					if (cfg.bin_polation == SGZ_BIN_NONE) { tapL(t.first[x]); if (!single) tapR(std::ptrdiff_t(N) - t.first[x]); }
					else if (cfg.bin_polation == SGZ_BIN_LINEAR)
					{
						const std::ptrdiff_t il = static_cast<std::ptrdiff_t>(std::floor(t.posL[x]));
						const std::ptrdiff_t ir = static_cast<std::ptrdiff_t>(std::floor(t.posR[x]));
						tapL(il); tapL(il + 1);
						if (!single) { tapR(ir); tapR(ir + 1); }
					}
					else
					{
						for (int k = 0; k < kLanczosTaps; ++k)
						{
							tapL(t.baseL[x] + k);
							if (!single) tapR(t.baseR[x] + k);
						}
					}
				}
				else
				{
					// scanned bins must stay inside [0, N/2] ([0, N] for Complex)
					const int32_t top = cplx ? int32_t(N) : int32_t(M);
					if (t.first[x] < 0 || t.first[x] + t.count[x] - 1 > top) ok = false;
					if (t.fallback[x] < 0 || t.fallback[x] > top) ok = false;
				}
			}
			t.linLimit = cplx ? N + 1 : std::min<uint32_t>(L, M + 1);
			// the fused kernel packs the bin map into 32 bits and derives fallback = first + count - 1 and the
			// channel-1 Lanczos base = first - 4; verify those identities and the field widths
			uint32_t safeFrom = 0;
			for (uint32_t x = 0; x < P; ++x)
			{
				if (t.first[x] < 0 || t.first[x] > 0xFFFF || t.count[x] < 0 || t.count[x] > 0x7FFF) ok = false;
				// (Phase starts its scan from bin 0 instead, TransformDSP.inl:789: the fused kernel does the same)
				if (t.mode[x] == 1 && t.fallback[x] != (phase ? 0 : t.first[x] + t.count[x] - 1)) ok = false;
				if (t.mode[x] == 0 && cfg.bin_polation == SGZ_BIN_LANCZOS)
				{
					if (t.baseL[x] != t.first[x] - (kLanczosSize - 1)) ok = false;
					const bool inL = t.baseL[x] >= 0 && t.baseL[x] + kLanczosTaps - 1 <= int32_t(N);
					const bool inR = single || (t.baseR[x] >= 0 && t.baseR[x] + kLanczosTaps - 1 <= int32_t(N));
					if (!(inL && inR)) safeFrom = x + 1;
				}
			}
			t.safeFrom = safeFrom;
			t.fusedOk = ok;
		}

		// ---- resonator (cpl::dsp::CComplexResonator is absent: SURVEY.md Appendix B decision) ----
		{
			const uint32_t V = static_cast<uint32_t>(2 * terms.size() - 1); // windowCoefficients(window).second
			t.resonatorVectors = V;
			t.resCoeffReal.assign(std::size_t(V) * P, 0); t.resCoeffImag.assign(std::size_t(V) * P, 0);
			t.resGain.assign(P, 0); t.resWindowTaps.assign(V, 0);
			const double fs = t.sampleRate;
			const double minWindow = 8, maxWindow = double(W);
			for (uint32_t i = 0; i < P; ++i)
			{
				const uint32_t k = (i + 1 >= P) ? (P >= 2 ? P - 2 : 0) : i;
				const double hz = P >= 2 ? std::abs(double(mf[k + 1]) - double(mf[k])) : fs;
				double len = hz > 0 ? fs / hz : maxWindow;
				if (!cfg.free_q) len = std::min(len, maxWindow);
				len = std::max(len, minWindow);
				const double bw = fs / len;
				const double r = std::exp(-kPi * bw / fs);
				t.resGain[i] = static_cast<float>(2.0 * (1.0 - r));
				for (uint32_t z = 0; z < V; ++z)
				{
					const double m = double(z) - double(V - 1) * 0.5;
					const double omega = 2.0 * kPi * (double(mf[i]) + m * bw) / fs;
					t.resCoeffReal[std::size_t(z) * P + i] = static_cast<float>(r * std::cos(omega));
					t.resCoeffImag[std::size_t(z) * P + i] = static_cast<float>(r * std::sin(omega));
				}
			}
			for (uint32_t z = 0; z < V; ++z)
			{
				const int m = int(z) - int(V - 1) / 2;
				const std::size_t am = static_cast<std::size_t>(m < 0 ? -m : m);
				float c = 0;
				if (am == 0) c = 1;
				else if (am < terms.size()) c = static_cast<float>(((am & 1) ? -1.0 : 1.0) * terms[am] / (2.0 * terms[0]));
				t.resWindowTaps[z] = c;
			}
		}
		return SGZ_OK;
	}
}
