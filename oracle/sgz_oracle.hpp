// sgz_oracle.hpp — CPU restatement of Signalizer's Spectrum transform path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under signalizer_b200/ may include, link
// or call this; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, as the checker / CPU baseline.
//
// PINNED BY THE REFERENCE'S SOURCE, cpl excepted: what is in-tree
// (Source/Spectrum/TransformDSP.inl, TransformConstant.h, TransformPair.h) is
// restated statement by statement below, with the same integer/float types at
// every conversion, and tests/test_oracle_ref.py holds this restatement to the
// same BITS as oracle/_ref - those three reference files compiled unmodified
// against stand-in cpl headers (oracle/build_ref.sh, oracle/ref_stubs/).
// What stays unpinned: the reference (jthorborg/signalizer @ 8286702) ships no
// tests, fixtures or golden vectors, and its numeric library `cpl`
// (bitbucket.org/Mayae/cpl @ fa8e21e1, a git submodule, which itself bundles
// pffft and DustFFT) is absent from the checkout.  The cpl primitives the path
// calls are restated from their published/textbook definitions as constrained
// by the call sites; each such decision is marked "cpl decision", and both this
// file and oracle/_ref run on them.
//
// Paths cited as `TransformDSP.inl:N` are relative to
// /root/reference/Source/Spectrum/.
#pragma once

#include <algorithm>
#include <array>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>
#include <sys/types.h>

namespace sgzo
{
	// CommonSignalizer.h:495-539
	enum class SpectrumChannels : int { Left, Right, Merge, Mid = Merge, Side, OffsetForMono = Side, Phase, Separate, MidSide, Complex, End };
	// SpectrumParameters.h:48-75
	enum LineGraphs { LineMain = 0, LineSecond, LineEnd };
	enum class BinInterpolation : int { None, Linear, Lanczos };
	enum class DisplayMode : int { LineGraph, ColourSpectrum };
	enum class TransformAlgorithm : int { FFT, RSNT };
	enum class ViewScaling : int { Linear, Logarithmic };
	// cpl decision: window family
	enum class WindowTypes : int { Rectangular, Hann, Hamming, Blackman, BlackmanHarris, Nuttall, FlatTop, End };
	enum class WindowShape : int { Symmetric, Periodic };

	// CommonSignalizer.h:541-591 (layout only)
	template<typename Scalar>
	struct UComplexFilter
	{
		Scalar a{ 0 }, b{ 0 }; // {real,imag} / {magnitude,phase} / {leftMagnitude,rightMagnitude}
	};

	namespace Math
	{
		// cpl decision: next power of two >= x
		inline std::size_t nextPow2Inc(std::size_t x)
		{
			std::size_t p = 1;
			while (p < x) p <<= 1;
			return p;
		}
		template<typename T> inline T confineTo(T v, T lo, T hi) { return v < lo ? lo : (v > hi ? hi : v); }
		// cpl decision: |z|^2 (assigned to a real and compared, TransformDSP.inl:625-631)
		template<typename T> inline T square(const std::complex<T>& z) { return z.real() * z.real() + z.imag() * z.imag(); }
		inline double dbToFraction(double db) { return std::pow(10.0, db / 20.0); }
	}

	// ---------------------------------------------------------------------------------------
	// cpl primitives (absent sources; definitions adopted per SURVEY.md Appendix B)
	// ---------------------------------------------------------------------------------------
	namespace dsp
	{
		// cpl decision: cosine-sum coefficients a_k, w[n] = sum_k (-1)^k a_k cos(2 pi k n / D)
		inline std::vector<double> windowCosineTerms(WindowTypes type)
		{
			switch (type)
			{
			case WindowTypes::Hann: return { 0.5, 0.5 };
			case WindowTypes::Hamming: return { 0.54, 0.46 };
			case WindowTypes::Blackman: return { 0.42, 0.5, 0.08 };
			case WindowTypes::BlackmanHarris: return { 0.35875, 0.48829, 0.14128, 0.01168 };
			case WindowTypes::Nuttall: return { 0.355768, 0.487396, 0.144232, 0.012604 };
			case WindowTypes::FlatTop: return { 0.21557895, 0.41663158, 0.277263158, 0.083578947, 0.006947368 };
			default: return { 1.0 };
			}
		}

		// cpl::dsp::windowCoefficients<T>(type).second: number of frequency-domain taps of the
		// window = number of resonator "vectors" (Spectrum.cpp:593).
		inline std::size_t windowCoefficientCount(WindowTypes type) { return 2 * windowCosineTerms(type).size() - 1; }

		// cpl::ParameterWindowDesignValue::generateWindow<T>(kernel, W) -> windowKernelScale.
		// cpl decision: scale = reciprocal coherent gain W / sum(w), so that
		// invSize = scale / (W / 2) maps a bin-centred sine of amplitude A to A
		// ("this will make scaling correct regardless of amount of zero-padding",
		// TransformDSP.inl:537-540).
		template<typename T>
		inline T generateWindow(T* kernel, std::size_t W, WindowTypes type, WindowShape shape)
		{
			const auto terms = windowCosineTerms(type);
			const double D = shape == WindowShape::Periodic ? double(W) : double(W > 1 ? W - 1 : 1);
			const double twoPi = 6.283185307179586476925286766559;
			double sum = 0;
			for (std::size_t n = 0; n < W; ++n)
			{
				double w = 0, sign = 1;
				for (std::size_t k = 0; k < terms.size(); ++k)
				{
					w += sign * terms[k] * std::cos(twoPi * double(k) * double(n) / D);
					sign = -sign;
				}
				kernel[n] = static_cast<T>(w);
				sum += static_cast<double>(kernel[n]);
			}
			return static_cast<T>(sum > 0 ? double(W) / sum : 1.0);
		}

		inline double sinc(double x)
		{
			if (x == 0) return 1.0;
			const double px = 3.1415926535897932384626433832795 * x;
			return std::sin(px) / px;
		}

		// cpl::dsp::linearFilter<T>(v, x): v[floor x](1 - frac) + v[floor x + 1] frac,
		// out-of-range taps contribute 0.  cpl decision: weights in the scalar type of T.
		template<typename S>
		inline std::complex<S> linearFilter(const std::complex<S>* v, std::ptrdiff_t size, S x)
		{
			const S fl = std::floor(x);
			const std::ptrdiff_t i = static_cast<std::ptrdiff_t>(fl);
			const S frac = x - fl;
			std::complex<S> a = (i >= 0 && i < size) ? v[i] : std::complex<S>(0);
			std::complex<S> b = (i + 1 >= 0 && i + 1 < size) ? v[i + 1] : std::complex<S>(0);
			return a * (S(1) - frac) + b * frac;
		}

		// Lanczos kernel weight for distance d (in bins/samples), lobe count a.
		// cpl decision: evaluated in double, rounded once to the scalar type.
		template<typename S>
		inline S lanczosWeight(double d, int a)
		{
			if (d <= -a || d >= a) return S(0);
			return static_cast<S>(sinc(d) * sinc(d / a));
		}

		// cpl::dsp::lanczosFilter<T, true>(v, x, a):
		// sum_{i = floor(x) - a + 1}^{floor(x) + a} v[i] L(x - i); out-of-range taps 0.
		// cpl decision: position arithmetic in the scalar type (x arrives as T from
		// `mappedFrequencies[x] * freqToBin`, or `N - that`), accumulation in the scalar type in
		// ascending tap order.
		template<typename S>
		inline std::complex<S> lanczosFilter(const std::complex<S>* v, std::ptrdiff_t size, S x, int a)
		{
			const S fl = std::floor(x);
			const std::ptrdiff_t base = static_cast<std::ptrdiff_t>(fl);
			std::complex<S> acc(0);
			for (std::ptrdiff_t i = base - a + 1; i <= base + a; ++i)
			{
				if (i < 0 || i >= size) continue;
				const S w = lanczosWeight<S>(static_cast<double>(x) - static_cast<double>(i), a);
				acc += v[i] * w;
			}
			return acc;
		}

		// real-vector variant used by the oscilloscope, lanczosFilter<double>(ptr, size, x, a)
		template<typename S, typename V>
		inline S lanczosFilterReal(const V* v, std::ptrdiff_t size, S x, int a)
		{
			const S fl = std::floor(x);
			const std::ptrdiff_t base = static_cast<std::ptrdiff_t>(fl);
			S acc(0);
			for (std::ptrdiff_t i = base - a + 1; i <= base + a; ++i)
			{
				if (i < 0 || i >= size) continue;
				acc += static_cast<S>(v[i]) * lanczosWeight<S>(static_cast<double>(x) - static_cast<double>(i), a);
			}
			return acc;
		}

		// cpl::dsp::separateTransformsIPL(Z): two-for-one split, in place.
		// After it Re Z[0] = ch1 DC, Im Z[0] = ch2 DC (TransformDSP.inl:861-862); channel 2
		// bin k is read at Z[N - k] (:893, :933); Z[N/2] is left packed.
		template<typename S>
		inline void separateTransformsIPL(std::complex<S>* Z, std::size_t N)
		{
			for (std::size_t k = 1; k < N / 2; ++k)
			{
				const std::complex<S> zk = Z[k], zn = std::conj(Z[N - k]);
				const std::complex<S> A = (zk + zn) * S(0.5);
				const std::complex<S> D = (zk - zn) * S(0.5);
				// B = D / i = (Im D, -Re D)
				Z[k] = A;
				Z[N - k] = std::complex<S>(D.imag(), -D.real());
			}
		}

		// cpl::dsp::UniFFT<std::complex<T>>{n}.forward(in, out, work): unnormalised forward DFT,
		// natural order, in == out allowed (TransformDSP.inl:494-498).  cpl decision (backend is
		// pffft for float per the licence list): X[k] = sum x[n] exp(-2 pi i k n / N).  This is a
		// Stockham radix-4 (+ one radix-2) autosort FFT with a double-derived twiddle table.
		template<typename S>
		class UniFFT
		{
		public:
			UniFFT() = default;
			explicit UniFFT(std::size_t n) : size(n), twiddle(n)
			{
				const double twoPi = 6.283185307179586476925286766559;
				for (std::size_t k = 0; k < n; ++k)
				{
					const double a = -twoPi * double(k) / double(n);
					twiddle[k] = std::complex<S>(static_cast<S>(std::cos(a)), static_cast<S>(std::sin(a)));
				}
			}

			void forward(const std::complex<S>* in, std::complex<S>* out, std::complex<S>* work) const
			{
				typedef std::complex<S> C;
				const std::size_t N = size;
				if (N == 0) return;
				if (N == 1) { out[0] = in[0]; return; }
				// count passes to decide the ping-pong parity so the result lands in `out`
				std::size_t passes = 0;
				for (std::size_t n = N; n > 1;) { if (n % 4 == 0) n /= 4; else n /= 2; ++passes; }
				C* x; C* y;
				if (passes % 2 == 0)
				{
					if (in != out) std::memcpy(out, in, N * sizeof(C));
					x = out; y = work;
				}
				else
				{
					std::memcpy(work, in, N * sizeof(C));
					x = work; y = out;
				}
				std::size_t n = N, s = 1;
				while (n > 1)
				{
					if (n % 4 == 0)
					{
						const std::size_t n1 = n / 4;
						for (std::size_t p = 0; p < n1; ++p)
						{
							const C w1 = twiddle[p * s], w2 = twiddle[2 * p * s], w3 = twiddle[3 * p * s];
							const C* x0 = x + s * p; const C* x1 = x + s * (p + n1);
							const C* x2 = x + s * (p + 2 * n1); const C* x3 = x + s * (p + 3 * n1);
							C* y0 = y + s * (4 * p); C* y1 = y0 + s; C* y2 = y1 + s; C* y3 = y2 + s;
							for (std::size_t q = 0; q < s; ++q)
							{
								const S ar = x0[q].real(), ai = x0[q].imag(), br = x1[q].real(), bi = x1[q].imag();
								const S cr = x2[q].real(), ci = x2[q].imag(), dr = x3[q].real(), di = x3[q].imag();
								const S apcr = ar + cr, apci = ai + ci, amcr = ar - cr, amci = ai - ci;
								const S bpdr = br + dr, bpdi = bi + di;
								// -i (b - d)
								const S jr = (bi - di), ji = -(br - dr);
								y0[q] = C(apcr + bpdr, apci + bpdi);
								const S t1r = amcr + jr, t1i = amci + ji;
								const S t2r = apcr - bpdr, t2i = apci - bpdi;
								const S t3r = amcr - jr, t3i = amci - ji;
								y1[q] = C(t1r * w1.real() - t1i * w1.imag(), t1r * w1.imag() + t1i * w1.real());
								y2[q] = C(t2r * w2.real() - t2i * w2.imag(), t2r * w2.imag() + t2i * w2.real());
								y3[q] = C(t3r * w3.real() - t3i * w3.imag(), t3r * w3.imag() + t3i * w3.real());
							}
						}
						n /= 4; s *= 4;
					}
					else
					{
						const std::size_t n1 = n / 2;
						for (std::size_t p = 0; p < n1; ++p)
						{
							const C w1 = twiddle[p * s];
							const C* x0 = x + s * p; const C* x1 = x + s * (p + n1);
							C* y0 = y + s * (2 * p); C* y1 = y0 + s;
							for (std::size_t q = 0; q < s; ++q)
							{
								const S ar = x0[q].real(), ai = x0[q].imag(), br = x1[q].real(), bi = x1[q].imag();
								y0[q] = C(ar + br, ai + bi);
								const S tr = ar - br, ti = ai - bi;
								y1[q] = C(tr * w1.real() - ti * w1.imag(), tr * w1.imag() + ti * w1.real());
							}
						}
						n /= 2; s *= 2;
					}
					std::swap(x, y);
				}
			}

			std::size_t size{ 0 };
		private:
			std::vector<std::complex<S>> twiddle;
		};

		// cpl::dsp::goertzel(buf, N, omega) -> complex<double> (OscilloscopeDSP.inl:277-285).
		// cpl decision: sum x[n] exp(-i omega n)
		template<typename V>
		inline std::complex<double> goertzel(const V* x, std::size_t N, double omega)
		{
			// second-order recurrence, finalised to the DFT-at-omega value
			const double c = 2 * std::cos(omega);
			double s1 = 0, s2 = 0;
			for (std::size_t n = 0; n < N; ++n)
			{
				const double s0 = static_cast<double>(x[n]) + c * s1 - s2;
				s2 = s1; s1 = s0;
			}
			// y = s[N-1] - e^{-i omega} s[N-2] = sum x[n] e^{i omega (N-1-n)}; rotate back by (N-1) omega
			const std::complex<double> y = s1 - std::complex<double>(std::cos(omega), -std::sin(omega)) * s2;
			const double back = -omega * double(N - 1);
			return y * std::complex<double>(std::cos(back), std::sin(back));
		}
	}

	// cpl::CPeakFilter<T>: only `.pole` is read on the path (TransformDSP.inl:1336).
	// cpl decision for setSampleRate / setDecayAsFraction(fraction, time): the pole that decays to
	// `fraction` after `time` seconds of frames at `rate` frames per second.
	template<typename T>
	struct CPeakFilter
	{
		T pole{ 0 };
		double sampleRate{ 1 };
		void setSampleRate(T rate) { sampleRate = rate; }
		void setDecayAsFraction(double fraction, double time)
		{
			pole = static_cast<T>(std::pow(fraction, 1.0 / (time * sampleRate)));
		}
	};

	// ---------------------------------------------------------------------------------------
	// cpl::AudioStream::AudioBufferView: a 2-segment (ring) view of one channel's history,
	// oldest to newest (consumed at TransformDSP.inl:65-88).
	// ---------------------------------------------------------------------------------------
	struct AudioBufferView
	{
		static constexpr std::size_t bufferIndices = 2;
		const float* seg[2]{ nullptr, nullptr };
		std::size_t len[2]{ 0, 0 };
		std::size_t size() const { return len[0] + len[1]; }
		const float* getItIndex(std::size_t i) const { return seg[i]; }
		std::size_t getItRange(std::size_t i) const { return len[i]; }
	};
	typedef std::array<AudioBufferView, 2> AudioPair;

	// ---------------------------------------------------------------------------------------
	// cpl::dsp::CComplexResonator<T, 2> (absent).  cpl decision, SURVEY.md Appendix B:
	// per filter i and vector z a complex one-pole s <- p s + x, p = r exp(i w);
	// bandwidth = spacing of neighbouring mapped frequencies, window length L = fs / bw samples
	// limited to [minWindow, maxWindow] unless freeQ; r = exp(-pi bw / fs); vectors sit at
	// f_i + m bw, m = z - (V-1)/2; the windowed snapshot combines them with the window's
	// frequency-domain taps c_m = (-1)^|m| a_|m| / (2 a_0) (c_0 = 1) and normalises by 2 (1 - r)
	// so a resonant sine of amplitude A reads A, like the FFT path.
	// ---------------------------------------------------------------------------------------
	template<typename T>
	struct ResonatorConstant
	{
		std::size_t numFilters{ 0 }, numVectors{ 0 };
		// [vector][filter]
		std::vector<T> coeffReal, coeffImag;
		std::vector<T> gain; // 2 (1 - r_i)
		std::size_t getNumFilters() const { return numFilters; }

		template<class Vector>
		void mapSystemHz(const Vector& mappedHz, std::size_t vectorSize, std::size_t vectors, T sampleRate, bool freeQ, std::size_t minWindow, std::size_t maxWindow)
		{
			numFilters = vectorSize; numVectors = vectors;
			coeffReal.assign(vectors * vectorSize, 0); coeffImag.assign(vectors * vectorSize, 0);
			gain.assign(vectorSize, 0);
			const double fs = sampleRate;
			const double pi = 3.1415926535897932384626433832795;
			for (std::size_t i = 0; i < vectorSize; ++i)
			{
				const std::size_t k = (i + 1 >= vectorSize) ? (vectorSize >= 2 ? vectorSize - 2 : 0) : i;
				double hz = vectorSize >= 2 ? std::abs(double(mappedHz[k + 1]) - double(mappedHz[k])) : fs;
				double L = hz > 0 ? fs / hz : double(maxWindow);
				if (!freeQ) L = std::min(L, double(maxWindow));
				L = std::max(L, double(minWindow));
				const double bw = fs / L;
				const double r = std::exp(-pi * bw / fs);
				gain[i] = static_cast<T>(2.0 * (1.0 - r));
				for (std::size_t z = 0; z < vectors; ++z)
				{
					const double m = double(z) - double(vectors - 1) * 0.5;
					const double omega = 2.0 * pi * (double(mappedHz[i]) + m * bw) / fs;
					coeffReal[z * vectorSize + i] = static_cast<T>(r * std::cos(omega));
					coeffImag[z * vectorSize + i] = static_cast<T>(r * std::sin(omega));
				}
			}
		}
	};

	template<typename T>
	struct ComplexResonator
	{
		// state [channel][vector][filter]
		std::vector<T> real, imag;
		std::size_t filters{ 0 }, vectors{ 0 };

		void ensure(const ResonatorConstant<T>& c)
		{
			if (filters != c.numFilters || vectors != c.numVectors)
			{
				filters = c.numFilters; vectors = c.numVectors;
				real.assign(2 * filters * vectors, 0); imag.assign(2 * filters * vectors, 0);
			}
		}
		void resetState() { std::fill(real.begin(), real.end(), T(0)); std::fill(imag.begin(), imag.end(), T(0)); }

		// resonateReal<V>(constant, work, channels, numSamples): TransformDSP.inl:1254-1291
		void resonateReal(const ResonatorConstant<T>& c, const T* const* work, std::size_t channels, std::size_t numSamples)
		{
			ensure(c);
			for (std::size_t ch = 0; ch < channels; ++ch)
			{
				for (std::size_t z = 0; z < vectors; ++z)
				{
					T* sr = real.data() + (ch * vectors + z) * filters;
					T* si = imag.data() + (ch * vectors + z) * filters;
					const T* pr = c.coeffReal.data() + z * filters;
					const T* pi = c.coeffImag.data() + z * filters;
					for (std::size_t i = 0; i < filters; ++i)
					{
						T re = sr[i], im = si[i];
						const T cr = pr[i], ci = pi[i];
						for (std::size_t n = 0; n < numSamples; ++n)
						{
							const T nr = re * cr - im * ci + work[ch][n];
							const T ni = re * ci + im * cr;
							re = nr; im = ni;
						}
						sr[i] = re; si[i] = im;
					}
				}
			}
		}

		// getWholeWindowedState<ISA>(constant, windowType, out, outChannels, numFilters):
		// out = interleaved complex, channel c at offset numFilters (TransformDSP.inl:1106-1107,1160)
		void getWholeWindowedState(const ResonatorConstant<T>& c, WindowTypes window, T* out, std::size_t outChannels, std::size_t numFilters)
		{
			ensure(c);
			const auto terms = dsp::windowCosineTerms(window);
			for (std::size_t ch = 0; ch < outChannels; ++ch)
			{
				for (std::size_t i = 0; i < numFilters; ++i)
				{
					T accr = 0, acci = 0;
					for (std::size_t z = 0; z < vectors; ++z)
					{
						const std::ptrdiff_t m = std::ptrdiff_t(z) - std::ptrdiff_t(vectors - 1) / 2;
						const std::size_t am = static_cast<std::size_t>(m < 0 ? -m : m);
						T cz = 0;
						if (am == 0) cz = 1;
						else if (am < terms.size()) cz = static_cast<T>(((am & 1) ? -1.0 : 1.0) * terms[am] / (2.0 * terms[0]));
						accr += cz * real[(ch * vectors + z) * filters + i];
						acci += cz * imag[(ch * vectors + z) * filters + i];
					}
					out[2 * (ch * numFilters + i)] = accr * c.gain[i];
					out[2 * (ch * numFilters + i) + 1] = acci * c.gain[i];
				}
			}
		}
	};

	// ---------------------------------------------------------------------------------------
	// TransformConstant<T> — TransformConstant.h:47-241
	// ---------------------------------------------------------------------------------------
	template<typename T>
	class TransformConstant
	{
	public:
		// TransformConstant.h:81-92
		void setStorage(std::size_t elements, std::size_t effectiveWindowSize, std::size_t& outputTransformSize)
		{
			windowSize = effectiveWindowSize;
			const auto newTransformSize = std::max<std::size_t>(32, Math::nextPow2Inc(windowSize));
			axisPoints = elements;
			transformSize = outputTransformSize = newTransformSize;
			windowKernel.resize(transformSize);
			mappedFrequencies.resize(axisPoints);
			slopeMap.resize(axisPoints);
			fft = dsp::UniFFT<T>(transformSize);
		}

		T mapFrequency(std::size_t axisPoint) const { return mappedFrequencies.at(axisPoint); }

		// TransformConstant.h:104-107
		void regenerateWindowKernel(WindowTypes type, WindowShape shape)
		{
			dspWindow = type;
			windowKernelScale = dsp::generateWindow<T>(windowKernel.data(), windowSize, type, shape);
		}

		// TransformConstant.h:109-118
		void generateSlopeMap(double fa, double fb)
		{
			const auto a = static_cast<T>(fa);
			const auto b = static_cast<T>(fb);
			for (std::size_t i = 0; i < axisPoints; ++i)
				slopeMap[i] = b * static_cast<T>(std::pow(mappedFrequencies[i], a));
		}

		// TransformConstant.h:120-123
		void remapResonator(bool shouldHaveFreeQ, std::size_t numVectors)
		{
			resonator.mapSystemHz(mappedFrequencies, mappedFrequencies.size(), numVectors, sampleRate, shouldHaveFreeQ, 8, windowSize);
		}

		// TransformConstant.h:125-180
		void remapFrequencies(double viewLeft, double viewRight, ViewScaling scaling, double minFreq)
		{
			double viewSize = viewRight - viewLeft; // Bounds<double>::dist()
			switch (scaling)
			{
			case ViewScaling::Linear:
			{
				double halfSampleRate = sampleRate * 0.5;
				double complexFactor = configuration == SpectrumChannels::Complex ? 2.0 : 1.0;
				double freqPerPixel = halfSampleRate / (axisPoints - 1);
				for (std::size_t i = 0; i < axisPoints; ++i)
					mappedFrequencies[i] = static_cast<T>(complexFactor * viewLeft * halfSampleRate + complexFactor * viewSize * i * freqPerPixel);
				break;
			}
			case ViewScaling::Logarithmic:
			{
				double sampleSize = double(axisPoints - 1);
				double end = sampleRate / 2;
				if (configuration != SpectrumChannels::Complex)
				{
					for (std::size_t i = 0; i < axisPoints; ++i)
						mappedFrequencies[i] = static_cast<T>(minFreq * std::pow(end / minFreq, viewLeft + viewSize * (i / sampleSize)));
				}
				else
				{
					for (std::size_t i = 0; i < axisPoints; ++i)
					{
						auto arg = viewLeft + viewSize * i / sampleSize;
						if (arg < 0.5)
						{
							mappedFrequencies[i] = static_cast<T>(minFreq * std::pow(end / minFreq, arg * 2));
						}
						else
						{
							arg -= 0.5;
							auto power = minFreq * std::pow(end / minFreq, 1.0 - arg * 2);
							mappedFrequencies[i] = static_cast<T>(end + (end - power));
						}
					}
				}
				break;
			}
			}
		}

		// TransformConstant.h:183-186
		std::size_t getStateConfigurationChannels() const noexcept
		{
			return configuration > SpectrumChannels::OffsetForMono ? 2 : 1;
		}

		ResonatorConstant<T> resonator;
		DisplayMode displayMode{};
		TransformAlgorithm algo{};
		BinInterpolation binPolation{};
		SpectrumChannels configuration{};
		WindowTypes dspWindow{};
		T sampleRate{};
		std::size_t axisPoints{}, transformSize{}, windowSize{};
		std::vector<T> mappedFrequencies;
		std::vector<T> windowKernel;
		T windowKernelScale{};
		std::size_t sampleBufferSize{ 200 };
		dsp::UniFFT<T> fft;
		std::array<CPeakFilter<T>, LineEnd> filter;
		double lowDBs{}, highDBs{}, clipDB{};
		std::vector<T> slopeMap;
	};

	// ---------------------------------------------------------------------------------------
	// TransformPair<T> — TransformPair.h:50-246, bodies TransformDSP.inl:38-1448
	// ---------------------------------------------------------------------------------------
	template<typename T>
	class TransformPair
	{
	public:
		typedef float AFloat;
		typedef UComplexFilter<AFloat> UComplex;
		typedef std::vector<UComplex> ComplexSpectrumFrame;
		typedef TransformConstant<T> Constant;
		typedef std::ptrdiff_t ssize_t_;
		typedef int fint_t;   // cpl::Types::fint_t
		typedef int fsint_t;  // cpl::Types::fsint_t

		struct LineGraphDesc
		{
			const std::vector<UComplex>& getResults(std::size_t size) const { results.resize(size); return results; }
			void resize(std::size_t n) { states.resize(n); results.resize(n); }
			void zero()
			{
				std::fill(states.begin(), states.end(), UComplex{});
				std::fill(results.begin(), results.end(), UComplex{});
			}
			std::vector<UComplex> states;
			mutable std::vector<UComplex> results;
		};

		std::vector<ComplexSpectrumFrame> sfbuf;
		std::array<LineGraphDesc, LineEnd> lineGraphs;
		std::size_t processedSamplesSinceLastFrame{};

		void clearLineGraphStates() { for (auto& g : lineGraphs) g.zero(); }
		void clearAudioState()
		{
			clearLineGraphStates();
			std::fill(workingMemory.begin(), workingMemory.end(), std::complex<T>(0));
			std::fill(audioMemory.begin(), audioMemory.end(), std::complex<T>(0));
			cresonator.resetState();
		}

		const std::complex<T>* getRawFFT(const Constant& constant) { return getAudioMemory(constant.transformSize); }
		// TransformDSP.inl:1444-1447
		const T* getTransformResult(const Constant& constant, std::size_t* size = nullptr)
		{
			const std::size_t n = constant.axisPoints * constant.getStateConfigurationChannels() * 2;
			if (size) *size = n;
			return getWorkT(n);
		}

		// TransformDSP.inl:39-231
		bool prepareTransform(const Constant& constant, const AudioPair& views)
		{
			if (views[0].size() != views[1].size() || views[0].size() < constant.windowSize)
				return false;

			std::size_t offset = views[0].size() - constant.windowSize;

			switch (constant.algo)
			{
			case TransformAlgorithm::FFT:
			{
				auto buffer = getAudioMemory(constant.transformSize);
				std::size_t channel = 1;
				std::size_t i = 0;

				switch (constant.configuration)
				{
				case SpectrumChannels::Left:
					channel = 0;
					[[fallthrough]];
				case SpectrumChannels::Right:
				{
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[channel].getItRange(indice);
						auto it = views[channel].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; it += offset;
							while (range--) { buffer[i] = *it++ * constant.windowKernel[i]; i++; }
							offset = 0;
						}
						else offset -= range;
					}
					break;
				}
				case SpectrumChannels::Merge:
				{
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[0].getItRange(indice);
						auto left = views[0].getItIndex(indice);
						auto right = views[1].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; left += offset; right += offset;
							while (range--) { buffer[i] = (*left++ + *right++) * constant.windowKernel[i] * 0.5f; i++; }
							offset = 0;
						}
						else offset -= range;
					}
					break;
				}
				case SpectrumChannels::Side:
				{
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[0].getItRange(indice);
						auto left = views[0].getItIndex(indice);
						auto right = views[1].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; left += offset; right += offset;
							while (range--) { buffer[i] = (*left++ - *right++) * constant.windowKernel[i] * 0.5f; i++; }
							offset = 0;
						}
						else offset -= range;
					}
					break;
				}
				case SpectrumChannels::MidSide:
				{
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[0].getItRange(indice);
						auto left = views[0].getItIndex(indice);
						auto right = views[1].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; left += offset; right += offset;
							while (range--)
							{
								buffer[i] = std::complex<T>(
									(*left + *right) * constant.windowKernel[i] * 0.5f,
									(*left - *right) * constant.windowKernel[i] * 0.5f);
								left++; right++; i++;
							}
							offset = 0;
						}
						else offset -= range;
					}
					break;
				}
				case SpectrumChannels::Phase:
				case SpectrumChannels::Separate:
				case SpectrumChannels::Complex:
				{
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[0].getItRange(indice);
						auto left = views[0].getItIndex(indice);
						auto right = views[1].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; left += offset; right += offset;
							while (range--)
							{
								buffer[i] = std::complex<T>{ *left++ * constant.windowKernel[i], *right++ * constant.windowKernel[i] };
								i++;
							}
							offset = 0;
						}
						else offset -= range;
					}
					break;
				}
				default: break;
				}
				for (std::size_t pad = i; pad < constant.transformSize; ++pad)
					buffer[pad] = (T)0;
				break;
			}
			default: break;
			}
			return true;
		}

		// TransformDSP.inl:234-484.  Restated literally, including the fact that in every
		// configuration except Left/Right the `left`/`right` iterators are NOT advanced by
		// `offset` (only the unused `it` is), so the history part is taken from the start of
		// each segment.
		bool prepareTransform(const Constant& constant, const AudioPair& views, std::array<const AFloat*, 2> preliminaryAudio, std::size_t numSamples)
		{
			if (views[0].size() != views[1].size() || views[0].size() < constant.windowSize)
				return false;

			std::size_t extraDiscardedSamples = views[0].size() - constant.windowSize;

			switch (constant.algo)
			{
			case TransformAlgorithm::FFT:
			{
				auto buffer = getAudioMemory(constant.transformSize);
				std::size_t channel = 1;
				std::size_t i = 0;
				std::size_t stop = std::min(numSamples, constant.windowSize);
				std::size_t offset = stop + extraDiscardedSamples;
				auto sizeToStopAt = constant.windowSize - offset;

				switch (constant.configuration)
				{
				case SpectrumChannels::Left:
					channel = 0;
					[[fallthrough]];
				case SpectrumChannels::Right:
				{
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[channel].getItRange(indice);
						auto it = views[channel].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; it += offset;
							while (range-- && i < sizeToStopAt) { buffer[i] = *it++ * constant.windowKernel[i]; i++; }
							offset = 0;
						}
						else offset -= range;
					}
					for (std::size_t k = 0; k < stop; ++i, k++)
						buffer[i] = preliminaryAudio[channel][k] * constant.windowKernel[i];
					break;
				}
				case SpectrumChannels::Merge:
				case SpectrumChannels::Side:
				case SpectrumChannels::MidSide:
				case SpectrumChannels::Phase:
				case SpectrumChannels::Separate:
				case SpectrumChannels::Complex:
				{
					const auto conf = constant.configuration;
					auto combine = [&](AFloat l, AFloat r, std::size_t idx) -> std::complex<T>
					{
						const T w = constant.windowKernel[idx];
						switch (conf)
						{
						case SpectrumChannels::Merge: return std::complex<T>((l + r) * w * (T)0.5, 0);
						case SpectrumChannels::Side: return std::complex<T>((l - r) * w * (T)0.5, 0);
						case SpectrumChannels::MidSide: return std::complex<T>((l + r) * w * (T)0.5, (l - r) * w * (T)0.5);
						default: return std::complex<T>(l * w, r * w);
						}
					};
					for (std::size_t indice = 0; indice < AudioBufferView::bufferIndices; ++indice)
					{
						std::size_t range = views[channel].getItRange(indice);
						auto left = views[0].getItIndex(indice);
						auto right = views[1].getItIndex(indice);
						if (range > offset)
						{
							range -= offset; // `it += offset` has no effect on left/right
							while (range-- && i < sizeToStopAt)
							{
								buffer[i] = combine(*left, *right, i);
								left++; right++; i++;
							}
							offset = 0;
						}
						else offset -= range;
					}
					for (std::size_t k = 0; k < stop; ++i, k++)
						buffer[i] = combine(preliminaryAudio[0][k], preliminaryAudio[1][k], i);
					break;
				}
				default: break;
				}
				for (std::size_t pad = i; pad < constant.transformSize; ++pad)
					buffer[pad] = 0;
				break;
			}
			default: break;
			}
			return true;
		}

		// TransformDSP.inl:487-502
		void doTransform(const Constant& constant)
		{
			switch (constant.algo)
			{
			case TransformAlgorithm::FFT:
				if (constant.transformSize > 0)
					constant.fft.forward(getAudioMemory(constant.transformSize), getAudioMemory(constant.transformSize), getWork(constant.transformSize));
				break;
			default: break;
			}
		}

		// TransformDSP.inl:506-1135
		void mapToLinearSpace(const Constant& constant)
		{
			switch (constant.algo)
			{
			case TransformAlgorithm::FFT:
			{
				const auto lanczosFilterSize = 5;
				ssize_t_ bin = 0, oldBin = 0, maxLBin, maxRBin = 0;
				fsint_t N = static_cast<fsint_t>(constant.transformSize);

				if (N < 3 || constant.sampleRate < 1)
					return;

				std::size_t numBins = N >> 1;
				auto const topFrequency = constant.sampleRate / 2;
				auto const freqToBin = static_cast<T>(numBins / topFrequency);

				std::complex<T> leftMax, rightMax;
				T maxLMag, maxRMag, newLMag, newRMag;

				// complex transform results, N + 1 size
				std::complex<T>* csf = getAudioMemory(constant.transformSize + 1);
				const ssize_t_ csfSize = static_cast<ssize_t_>(constant.transformSize + 1);
				// buffer for single results, numPoints * 2 size / complex results numPoints * 2 size (same memory)
				std::complex<T>* csp = getWork(constant.axisPoints * 2);
				T* wsp = reinterpret_cast<T*>(csp);

				auto const invSize = static_cast<T>(constant.windowKernelScale / (constant.windowSize * 0.5));
				const auto& mf = constant.mappedFrequencies;
				const fint_t axisPoints = static_cast<fint_t>(constant.axisPoints);

				switch (constant.configuration)
				{
				case SpectrumChannels::Left:
				case SpectrumChannels::Right:
				case SpectrumChannels::Merge:
				case SpectrumChannels::Side:
				{
					oldBin = static_cast<ssize_t_>(mf[0] * freqToBin);

					csf[0] *= (T)0.5;
					csf[N >> 1] *= (T)0.5;

					for (std::size_t i = 0; i < numBins; ++i)
						csf[i] = std::abs(csf[i]);

					double fftBandwidth = 1.0 / numBins;
					fint_t x = 0;
					switch (constant.binPolation)
					{
					case BinInterpolation::None:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							auto index = Math::confineTo<std::size_t>((std::size_t)(mf[x] * freqToBin + 0.5), 0, numBins);
							csp[x] = invSize * csf[index];
						}
						break;
					case BinInterpolation::Linear:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							csp[x] = invSize * dsp::linearFilter<T>(csf, csfSize, mf[x] * freqToBin);
						}
						break;
					case BinInterpolation::Lanczos:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							csp[x] = invSize * dsp::lanczosFilter<T>(csf, csfSize, mf[x] * freqToBin, lanczosFilterSize);
						}
						break;
					default:
						break;
					}

					oldBin = static_cast<ssize_t_>(mf[x] * freqToBin);

					for (; x < axisPoints; ++x)
					{
						maxLMag = maxRMag = newLMag = newRMag = 0;
						bin = static_cast<ssize_t_>(static_cast<std::size_t>(mf[x] * freqToBin));
						maxRBin = maxLBin = bin;

						auto diff = bin - oldBin;
						auto counter = diff ? 1 : 0;
						do
						{
							auto offset = oldBin + counter;
							newLMag = Math::square(csf[offset]);
							if (newLMag > maxLMag)
							{
								maxLBin = oldBin + counter;
								maxLMag = newLMag;
							}
							counter++;
							diff--;
						} while (diff > 0);

						csp[x] = invSize * csf[maxLBin];
						oldBin = bin;
					}
					break;
				}
				case SpectrumChannels::Phase:
				{
					dsp::separateTransformsIPL(csf, static_cast<std::size_t>(N));

					csf[N] = csf[0].imag() * (T)0.5;
					csf[0] = csf[0].real() * (T)0.5;
					csf[N >> 1] *= (T)0.5;
					csf[(N >> 1) - 1] *= (T)0.5;

					std::size_t bandWidthBreakingPoint = constant.axisPoints;

					double fftBandwidth = 1.0 / numBins;
					fint_t x = 0;
					switch (constant.binPolation)
					{
					case BinInterpolation::Linear:
					{
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
							{
								bandWidthBreakingPoint = x;
								break;
							}
							auto iLeft = dsp::linearFilter<T>(csf, csfSize, mf[x] * freqToBin);
							auto iRight = dsp::linearFilter<T>(csf, csfSize, N - (mf[x] * freqToBin));
							auto cancellation = invSize * std::sqrt(Math::square(iLeft + iRight));
							auto mid = invSize * (std::abs(iLeft) + std::abs(iRight));
							wsp[x * 2 + 1] = T(1) - (mid > 0 ? (cancellation / mid) : 0);
						}

						std::size_t normalizedPosition = 0U;
						const std::size_t linearFilterSize = 1;

						for (x = 0; static_cast<std::size_t>(x) < bandWidthBreakingPoint; ++x)
						{
							auto binPosition = mf[x] * freqToBin;
							while ((binPosition + linearFilterSize) > normalizedPosition && static_cast<std::size_t>(x) < bandWidthBreakingPoint - (linearFilterSize + 1))
							{
								csf[normalizedPosition] = std::abs(csf[normalizedPosition]);
								csf[N - normalizedPosition] = std::abs(csf[N - normalizedPosition]);
								normalizedPosition++;
							}
							auto iLeft = dsp::linearFilter<T>(csf, csfSize, binPosition);
							auto iRight = dsp::linearFilter<T>(csf, csfSize, N - binPosition);
							wsp[x * 2] = invSize * (std::abs(iLeft) + std::abs(iRight));
						}
						break;
					}
					case BinInterpolation::Lanczos:
					{
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
							{
								bandWidthBreakingPoint = x;
								break;
							}
							auto iLeft = dsp::lanczosFilter<T>(csf, csfSize, mf[x] * freqToBin, lanczosFilterSize);
							auto iRight = dsp::lanczosFilter<T>(csf, csfSize, N - (mf[x] * freqToBin), lanczosFilterSize);
							auto cancellation = invSize * std::sqrt(Math::square(iLeft + iRight));
							auto mid = invSize * (std::abs(iLeft) + std::abs(iRight));
							wsp[x * 2 + 1] = T(1) - (mid > 0 ? (cancellation / mid) : 0);
						}

						std::size_t normalizedPosition = 0U;

						for (x = 0; static_cast<std::size_t>(x) < bandWidthBreakingPoint; ++x)
						{
							auto binPosition = mf[x] * freqToBin;
							// `bandWidthBreakingPoint - (lanczosFilterSize + 1)`: size_t minus int, wraps when < 6
							while ((binPosition + lanczosFilterSize) > normalizedPosition && static_cast<std::size_t>(x) < bandWidthBreakingPoint - static_cast<std::size_t>(lanczosFilterSize + 1))
							{
								csf[normalizedPosition] = std::abs(csf[normalizedPosition]);
								csf[N - normalizedPosition] = std::abs(csf[N - normalizedPosition]);
								normalizedPosition++;
							}
							auto iLeft = dsp::lanczosFilter<T>(csf, csfSize, mf[x] * freqToBin, lanczosFilterSize);
							auto iRight = dsp::lanczosFilter<T>(csf, csfSize, N - (mf[x] * freqToBin), lanczosFilterSize);
							wsp[x * 2] = invSize * (std::abs(iLeft) + std::abs(iRight));
						}
						break;
					}
					default:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							auto index = Math::confineTo<std::size_t>((std::size_t)(mf[x] * freqToBin + 0.5), 0, numBins - 1);
							auto iLeft = csf[index];
							auto iRight = csf[N - index];
							auto cancellation = invSize * std::abs(iLeft + iRight);
							auto mid = invSize * (std::abs(iLeft) + std::abs(iRight));
							wsp[x * 2] = mid;
							wsp[x * 2 + 1] = T(1) - (mid > 0 ? (cancellation / mid) : 0);
						}
						break;
					}

					if (x < axisPoints)
						oldBin = static_cast<ssize_t_>(mf[x] * freqToBin);

					for (; x < axisPoints; ++x)
					{
						std::size_t maxBin = 0;
						T maxValue = 0, newMag = 0;
						bin = static_cast<ssize_t_>(static_cast<std::size_t>(mf[x] * freqToBin));

						auto diff = bin - oldBin;
						auto counter = diff ? 1 : 0;
						do
						{
							auto offset = oldBin + counter;
							newMag = std::max(Math::square(csf[offset]), Math::square(csf[N - offset]));
							if (newMag > maxValue)
							{
								maxValue = newMag;
								maxBin = oldBin + counter;
							}
							counter++;
							diff--;
						} while (diff > 0);

						leftMax = csf[maxBin];
						rightMax = csf[N - maxBin];

						auto interference = invSize * std::abs(leftMax + rightMax);
						auto mid = invSize * (std::abs(leftMax) + std::abs(rightMax));
						auto cancellation = interference / mid;
						wsp[x * 2] = mid;
						wsp[x * 2 + 1] = T(1) - (mid > 0 ? cancellation : 0);

						oldBin = bin;
					}
					break;
				}
				case SpectrumChannels::Separate:
				case SpectrumChannels::MidSide:
				{
					dsp::separateTransformsIPL(csf, static_cast<std::size_t>(N));

					csf[N] = csf[0].imag() * (T)0.5;
					csf[0] = csf[0].real() * (T)0.5;
					csf[N >> 1] *= (T)0.5;
					csf[(N >> 1) - 1] *= (T)0.5;

					for (fsint_t i = 1; i < N; ++i)
						csf[i] = std::abs(csf[i]);

					double fftBandwidth = 1.0 / numBins;
					fint_t x;
					switch (constant.binPolation)
					{
					case BinInterpolation::Linear:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							auto iLeft = dsp::linearFilter<T>(csf, csfSize, mf[x] * freqToBin);
							auto iRight = dsp::linearFilter<T>(csf, csfSize, N - (mf[x] * freqToBin));
							csp[x] = invSize * iLeft;
							csp[constant.axisPoints + x] = invSize * iRight;
						}
						break;
					case BinInterpolation::Lanczos:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							auto iLeft = dsp::lanczosFilter<T>(csf, csfSize, mf[x] * freqToBin, 5);
							auto iRight = dsp::lanczosFilter<T>(csf, csfSize, N - (mf[x] * freqToBin), lanczosFilterSize);
							csp[x] = invSize * iLeft;
							csp[constant.axisPoints + x] = invSize * iRight;
						}
						break;
					default:
						for (x = 0; x < axisPoints - 1; ++x)
						{
							double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
							if (bwForLine > fftBandwidth)
								break;
							auto index = Math::confineTo<std::size_t>((std::size_t)(mf[x] * freqToBin + 0.5), 0, numBins);
							csp[x] = invSize * csf[index];
							csp[constant.axisPoints + x] = invSize * csf[N - index];
						}
						break;
					}

					oldBin = static_cast<ssize_t_>(mf[x] * freqToBin);

					for (; x < axisPoints; ++x)
					{
						maxLMag = maxRMag = newLMag = newRMag = 0;
						bin = static_cast<ssize_t_>(static_cast<std::size_t>(mf[x] * freqToBin));
						maxRBin = maxLBin = bin;

						auto diff = bin - oldBin;
						auto counter = diff ? 1 : 0;
						do
						{
							auto offset = oldBin + counter;
							newLMag = Math::square(csf[offset]);
							newRMag = Math::square(csf[N - offset]);
							if (newLMag > maxLMag)
							{
								maxLBin = oldBin + counter;
								maxLMag = newLMag;
							}
							if (newRMag > maxRMag)
							{
								maxRBin = N - (oldBin + counter);
								maxRMag = newRMag;
							}
							counter++;
							diff--;
						} while (diff > 0);

						csp[x] = invSize * csf[maxLBin];
						csp[constant.axisPoints + x] = invSize * csf[maxRBin];
						oldBin = bin;
					}
					break;
				}
				case SpectrumChannels::Complex:
				{
					csf[0] *= (T)0.5;

					double fftBandwidth = 1.0 / (numBins * 2);
					fint_t x = 0;

					for (fsint_t i = 1; i < N; ++i)
						csf[i] = std::abs(csf[i]);

					while (x < axisPoints)
					{
						switch (constant.binPolation)
						{
						case BinInterpolation::Linear:
							for (; x < axisPoints; ++x)
							{
								if (x != axisPoints - 1)
								{
									double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
									if (bwForLine > fftBandwidth)
										break;
								}
								csp[x] = invSize * dsp::linearFilter<T>(csf, csfSize, mf[x] * freqToBin);
							}
							break;
						case BinInterpolation::Lanczos:
							for (; x < axisPoints; ++x)
							{
								if (x != axisPoints - 1)
								{
									double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
									if (bwForLine > fftBandwidth)
										break;
								}
								csp[x] = invSize * dsp::lanczosFilter<T>(csf, csfSize, mf[x] * freqToBin, lanczosFilterSize);
							}
							break;
						default:
							for (; x < axisPoints; ++x)
							{
								if (x != axisPoints - 1)
								{
									double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
									if (bwForLine > fftBandwidth)
										break;
								}
								auto index = Math::confineTo<std::size_t>((std::size_t)(mf[x] * freqToBin + 0.5), 0, N);
								csp[x] = invSize * csf[index];
							}
							break;
						}
						if (x != axisPoints)
							oldBin = static_cast<ssize_t_>(mf[x] * freqToBin);

						for (; x < axisPoints; ++x)
						{
							maxLMag = maxRMag = newLMag = newRMag = 0;
							bin = static_cast<ssize_t_>(static_cast<std::size_t>(mf[x] * freqToBin));
							maxRBin = maxLBin = bin;
							if (x != axisPoints - 1)
							{
								double bwForLine = (mf[x + 1] - mf[x]) / topFrequency;
								if (bwForLine < fftBandwidth)
									break;
							}

							auto diff = bin - oldBin;
							auto counter = diff ? 1 : 0;
							do
							{
								auto offset = oldBin + counter;
								newLMag = Math::square(csf[offset]);
								if (newLMag > maxLMag)
								{
									maxLBin = oldBin + counter;
									maxLMag = newLMag;
								}
								counter++;
								diff--;
							} while (diff > 0);

							csp[x] = invSize * csf[maxLBin];
							oldBin = bin;
						}
					}
					break;
				}
				default: break;
				}
				break;
			}
			case TransformAlgorithm::RSNT:
			{
				auto configurationChannels = constant.getStateConfigurationChannels();
				auto wsp = getWork(constant.resonator.getNumFilters() * configurationChannels);
				std::size_t filtersPerChannel = copyResonatorStateInto(constant, constant.dspWindow, wsp, configurationChannels) / configurationChannels;

				switch (constant.configuration)
				{
				case SpectrumChannels::Phase:
				{
					for (std::size_t x = 0; x < filtersPerChannel; ++x)
					{
						auto iLeft = wsp[x];
						auto iRight = wsp[x + filtersPerChannel];
						auto cancellation = std::sqrt(Math::square(iLeft + iRight));
						auto mid = std::abs(iLeft) + std::abs(iRight);
						wsp[x] = std::complex<T>(mid, T(1) - (mid > 0 ? (cancellation / mid) : 0));
					}
					break;
				}
				default: break;
				}
				break;
			}
			}
		}

		// TransformDSP.inl:1139-1148
		void addAudioFrame(const Constant& constant)
		{
			mapToLinearSpace(constant);
			postProcessStdTransform(constant);
			constexpr auto i = LineMain;
			auto frame = ComplexSpectrumFrame(lineGraphs[i].results.begin(), lineGraphs[i].results.begin() + constant.axisPoints);
			sfbuf.emplace_back(std::move(frame));
		}

		// TransformDSP.inl:1152-1161
		std::size_t copyResonatorStateInto(const Constant& constant, WindowTypes windowType, std::complex<T>* output, std::size_t outChannels)
		{
			auto numResFilters = constant.resonator.getNumFilters();
			cresonator.getWholeWindowedState(constant.resonator, windowType, reinterpret_cast<T*>(output), outChannels, numResFilters);
			return numResFilters << (outChannels - 1);
		}

		// TransformDSP.inl:1165-1211
		void audioEntryPoint(const Constant& constant, const AudioPair* views, std::array<const AFloat*, 2> buffer, std::size_t numSamples)
		{
			if (constant.displayMode == DisplayMode::ColourSpectrum)
			{
				std::int64_t n = static_cast<std::int64_t>(numSamples);
				std::size_t offset = 0;

				while (n > 0)
				{
					std::int64_t numRemainingSamples = processedSamplesSinceLastFrame > constant.sampleBufferSize ? 0 : constant.sampleBufferSize - processedSamplesSinceLastFrame;
					const auto availableSamples = numRemainingSamples + std::min(std::int64_t(0), n - numRemainingSamples);

					if (constant.algo == TransformAlgorithm::RSNT)
						resonatingDispatch(constant, { buffer[0] + offset, buffer[1] + offset }, static_cast<std::size_t>(availableSamples));

					processedSamplesSinceLastFrame += availableSamples;

					if (processedSamplesSinceLastFrame >= constant.sampleBufferSize)
					{
						bool transformReady = true;
						if (constant.algo == TransformAlgorithm::FFT)
						{
							// NOTE: passes `buffer` (not buffer + offset) exactly like the reference (:1192)
							if ((transformReady = prepareTransform(constant, *views, buffer, static_cast<std::size_t>(availableSamples))))
								doTransform(constant);
						}
						if (transformReady)
							addAudioFrame(constant);
						processedSamplesSinceLastFrame = 0;
					}

					offset += availableSamples;
					n -= availableSamples;
				}
			}
			else if (constant.algo == TransformAlgorithm::RSNT)
			{
				resonatingDispatch(constant, buffer, numSamples);
			}
		}

		// TransformDSP.inl:1215-1295
		void resonatingDispatch(const Constant& constant, std::array<const AFloat*, 2> buffer, std::size_t numSamples)
		{
			if (numSamples < 1)
				return;
			std::size_t channels = 1;
			T* work = getWorkT(numSamples * 2);
			auto one = [&](auto fn) { for (std::size_t i = 0; i < numSamples; ++i) work[i] = static_cast<T>(fn(buffer[0][i], buffer[1][i])); channels = 1; };
			auto two = [&](auto fn)
			{
				for (std::size_t i = 0; i < numSamples; ++i)
				{
					auto result = fn(buffer[0][i], buffer[1][i]);
					work[i] = static_cast<T>(result.first);
					work[i + numSamples] = static_cast<T>(result.second);
				}
				channels = 2;
			};
			switch (constant.configuration)
			{
			case SpectrumChannels::Right: one([](AFloat, AFloat right) { return right; }); break;
			case SpectrumChannels::Left: one([](AFloat left, AFloat) { return left; }); break;
			case SpectrumChannels::Mid: one([](AFloat left, AFloat right) { return left + right; }); break;
			case SpectrumChannels::Side: one([](AFloat left, AFloat right) { return left - right; }); break;
			case SpectrumChannels::MidSide: two([](AFloat left, AFloat right) { return std::pair<AFloat, AFloat>{ left - right, left + right }; }); break;
			case SpectrumChannels::Phase:
			case SpectrumChannels::Separate:
			case SpectrumChannels::Complex: two([](AFloat left, AFloat right) { return std::pair<AFloat, AFloat>{ left, right }; }); break;
			default: return;
			}
			const T* ptrs[2] = { work, work + numSamples };
			cresonator.resonateReal(constant.resonator, ptrs, channels, numSamples);
		}

		// TransformDSP.inl:1299-1435
		void mapAndTransformDFTFilters(const Constant& constant, const T* newVals, std::size_t size)
		{
			for (std::size_t k = 0; k < lineGraphs.size(); ++k)
				lineGraphs[k].resize(size);

			double lowerFraction = Math::dbToFraction(constant.lowDBs);
			double upperFraction = Math::dbToFraction(constant.highDBs);

			auto deltaYRecip = static_cast<T>(1.0 / std::log(upperFraction / lowerFraction));
			auto minFracRecip = static_cast<T>(1.0 / lowerFraction);
			T lowerClip = static_cast<T>(constant.clipDB);

			switch (constant.configuration)
			{
			case SpectrumChannels::Left:
			case SpectrumChannels::Merge:
			case SpectrumChannels::Right:
			case SpectrumChannels::Side:
			case SpectrumChannels::Complex:
			{
				for (fint_t i = 0; static_cast<std::size_t>(i) < size; ++i)
				{
					auto newReal = newVals[i * 2];
					auto newImag = newVals[i * 2 + 1];
					auto magnitude = std::sqrt(newReal * newReal + newImag * newImag);

					for (std::size_t k = 0; k < lineGraphs.size(); ++k)
					{
						lineGraphs[k].states[i].a *= constant.filter[k].pole;
						if (magnitude > lineGraphs[k].states[i].a)
							lineGraphs[k].states[i].a = static_cast<AFloat>(magnitude);

						auto deltaX = constant.slopeMap[i] * lineGraphs[k].states[i].a * minFracRecip;
						auto result = deltaX > 0 ? std::log(deltaX) * deltaYRecip : lowerClip;
						lineGraphs[k].results[i].a = static_cast<AFloat>(result);
						lineGraphs[k].results[i].b = 0;
					}
				}
				break;
			}
			case SpectrumChannels::Separate:
			case SpectrumChannels::MidSide:
			{
				for (fint_t i = 0; static_cast<std::size_t>(i) < size; ++i)
				{
					auto lreal = newVals[i * 2];
					auto rreal = newVals[i * 2 + size * 2];
					auto limag = newVals[i * 2 + 1];
					auto rimag = newVals[i * 2 + size * 2 + 1];
					auto lmag = std::sqrt(lreal * lreal + limag * limag);
					auto rmag = std::sqrt(rreal * rreal + rimag * rimag);

					for (std::size_t k = 0; k < lineGraphs.size(); ++k)
					{
						lineGraphs[k].states[i].a *= constant.filter[k].pole;
						lineGraphs[k].states[i].b *= constant.filter[k].pole;

						if (lmag > lineGraphs[k].states[i].a)
							lineGraphs[k].states[i].a = static_cast<AFloat>(lmag);
						if (rmag > lineGraphs[k].states[i].b)
							lineGraphs[k].states[i].b = static_cast<AFloat>(rmag);

						auto deltaLX = constant.slopeMap[i] * lineGraphs[k].states[i].a * minFracRecip;
						auto deltaRX = constant.slopeMap[i] * lineGraphs[k].states[i].b * minFracRecip;
						auto lResult = deltaLX > 0 ? std::log(deltaLX) * deltaYRecip : lowerClip;
						auto rResult = deltaRX > 0 ? std::log(deltaRX) * deltaYRecip : lowerClip;
						lineGraphs[k].results[i].a = static_cast<AFloat>(lResult);
						lineGraphs[k].results[i].b = static_cast<AFloat>(rResult);
					}
				}
				break;
			}
			case SpectrumChannels::Phase:
			{
				T phaseFilters[LineEnd];
				for (std::size_t k = 0; k < lineGraphs.size(); ++k)
					phaseFilters[k] = static_cast<T>(std::pow(constant.filter[k].pole, T(0.3)));

				for (fint_t i = 0; static_cast<std::size_t>(i) < size; ++i)
				{
					auto mag = newVals[i * 2];
					auto phase = newVals[i * 2 + 1];
					mag *= T(0.5);

					for (std::size_t k = 0; k < lineGraphs.size(); ++k)
					{
						lineGraphs[k].states[i].a *= constant.filter[k].pole;
						if (mag > lineGraphs[k].states[i].a)
							lineGraphs[k].states[i].a = static_cast<AFloat>(mag);
						phase *= mag;
						lineGraphs[k].states[i].b = static_cast<AFloat>(phase + phaseFilters[k] * (lineGraphs[k].states[i].b - phase));

						auto deltaX = constant.slopeMap[i] * lineGraphs[k].states[i].a * minFracRecip;
						auto deltaY = constant.slopeMap[i] * lineGraphs[k].states[i].b * minFracRecip;
						auto result = deltaX > 0 ? std::log(deltaX) * deltaYRecip : lowerClip;
						lineGraphs[k].results[i].a = static_cast<AFloat>(result);
						lineGraphs[k].results[i].b = static_cast<AFloat>(deltaY > 0 ? std::log(deltaY) * deltaYRecip : lowerClip);
					}
				}
				break;
			}
			default: break;
			}
		}

		// TransformDSP.inl:1438-1441
		void postProcessStdTransform(const Constant& constant)
		{
			mapAndTransformDFTFilters(constant, getTransformResult(constant), constant.axisPoints);
		}

		ComplexResonator<T> cresonator;

	private:
		// TransformPair.h:213-233: both are vectors of complex<T> that only ever grow
		std::complex<T>* getWork(std::size_t size)
		{
			if (workingMemory.size() < size) workingMemory.resize(size);
			return workingMemory.data();
		}
		T* getWorkT(std::size_t size)
		{
			const std::size_t c = (size + 1) / 2;
			if (workingMemory.size() < c) workingMemory.resize(c);
			return reinterpret_cast<T*>(workingMemory.data());
		}
		std::complex<T>* getAudioMemory(std::size_t size)
		{
			if (audioMemory.size() < size) audioMemory.resize(size);
			return audioMemory.data();
		}
		std::vector<std::complex<T>> workingMemory;
		std::vector<std::complex<T>> audioMemory;
	};
}
