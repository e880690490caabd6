// sgz_oracle_scope.hpp — CPU restatement of the secondary paths (Oscilloscope, Vectorscope).
//
// TEST INFRASTRUCTURE ONLY; PARITY UNPINNED (see sgz_oracle.hpp).  References are relative to
// /root/reference/Source/.
#pragma once

#include <algorithm>
#include <array>
#include <cmath>
#include <complex>
#include <cstdint>
#include <deque>
#include <vector>

#include "sgz_oracle.hpp"

namespace sgzo
{
	// ---------------------------------------------------------------------------------------
	// Vectorscope — Vectorscope/VectorscopeRendering.cpp:500-746 (polar), :444-497 (Lissajous)
	// ---------------------------------------------------------------------------------------
	struct Vertex { float x, y, z; };

	// cpl decision: cpl::simd::atan / sincos are SSE polynomial approximations of unknown form
	// (cpl is absent).  The pair is only ever used composed, (sx, cx) = sincos(atan(X / Y))
	// (:582-586), which equals (t, 1) / sqrt(1 + t^2) with t = X / Y exactly; evaluating that
	// with correctly rounded IEEE float operations makes CPU and GPU agree bit for bit, which the
	// histogram counts need.  cos(atan) >= 0 always, so the half-plane fold of the reference
	// (atan, not atan2) is preserved.
	inline void sincosOfAtanRatio(float X, float Y, float& sx, float& cx)
	{
		const float aX = std::fabs(X), aY = std::fabs(Y);
		if (aX == 0.0f && aY == 0.0f) { sx = 0.0f; cx = 1.0f; return; }
		if (aY >= aX)
		{
			const float t = X / Y;
			const float h = std::sqrt(1.0f + t * t);
			cx = 1.0f / h;
			sx = t / h;
		}
		else
		{
			const float u = Y / X;
			const float h = std::sqrt(1.0f + u * u);
			cx = std::fabs(u) / h;
			const float m = 1.0f / h;
			sx = (std::signbit(X) != std::signbit(Y)) ? -m : m;
		}
	}

	// one sample of drawPolarPlot (scalar route :608-632; the vector route :562-602 is the same math)
	inline Vertex polarVertex(float left, float right, float fade)
	{
		const float cosineRotation = -0.70710678118654752440f, sineRotation = 0.70710678118654752440f;
		const float length = std::max(std::fabs(left), std::fabs(right));
		const float vY = left * cosineRotation - right * sineRotation;
		const float vX = left * sineRotation + right * cosineRotation;
		float sx, cx;
		if (left == 0.0f && right == 0.0f) { sx = 0.0f; cx = 1.0f; } // angle = 0 (:627)
		else sincosOfAtanRatio(vX, vY, sx, cx);
		return Vertex{ sx * length, cx * length, fade };
	}

	struct VScopeParams
	{
		bool polar{ true }, scalePolar{ false };
		float gain{ 1 };                 // processor->envelopeGain * state.userGain (:512, :451)
		uint32_t width{ 512 }, height{ 512 };
	};

	// vertex after the model transform: gain (:517) and Conditional01To11HeightTransform (:286-300):
	// (g x, 2 g y - 1) when scalePolar.  z (fade) = i / numSamples - 1 (:528,:592; the reference
	// accumulates it per SIMD vector, which is ISA dependent — restated as one multiply).
	inline Vertex vscopeVertex(const VScopeParams& p, float left, float right, std::size_t i, std::size_t numSamples)
	{
		Vertex v;
		if (p.polar)
		{
			const float fadePerSample = 1.0f / float(numSamples);
			v = polarVertex(left, right, float(i) * fadePerSample - 1.0f);
		}
		else
		{
			const float sampleFade = 1.0f / float(std::max<int>(1, int(numSamples) - 1));
			v = Vertex{ right, left, float(i) * sampleFade - 1.0f }; // drawRectPlot :466
		}
		v.x = v.x * p.gain; v.y = v.y * p.gain;
		if (p.polar && p.scalePolar) v.y = v.y * 2.0f - 1.0f;
		return v;
	}

	// The point rasterisation the reference leaves to OpenGL (GL_POINTS), as a density histogram:
	// NDC [-1, 1) -> pixel floor((c + 1) * (size / 2)); points outside are clipped.
	inline void vscopeHistogram(const VScopeParams& p, const float* left, const float* right, std::size_t n, uint32_t* hist)
	{
		const float hw = 0.5f * float(p.width), hh = 0.5f * float(p.height);
		for (std::size_t i = 0; i < n; ++i)
		{
			const Vertex v = vscopeVertex(p, left[i], right[i], i, n);
			const float fx = std::floor((v.x + 1.0f) * hw), fy = std::floor((v.y + 1.0f) * hh);
			if (!(fx >= 0.0f && fx < float(p.width) && fy >= 0.0f && fy < float(p.height))) continue;
			hist[std::size_t(fy) * p.width + std::size_t(fx)]++;
		}
	}

	// Vectorscope/Vectorscope.cpp:268-377 — stereo meters (one-pole IIRs).  cos(2 atan(Y / X)) is
	// restated exactly as (1 - t^2) / (1 + t^2), t = Y / X (same cpl decision as above).
	struct VScopeFilters { float envelope[2]{ 0, 0 }, balance[2][2]{ { 0, 0 }, { 0, 0 } }, phase[2]{ 0, 0 }; };
	inline float cosTwiceAtan(float y, float x)
	{
		if (x == 0.0f && y == 0.0f) return 0.0f; // dummy angle pi/4 -> cos(pi/2) (:306-310)
		if (std::fabs(x) >= std::fabs(y)) { const float t = y / x; const float t2 = t * t; return (1.0f - t2) / (1.0f + t2); }
		const float u = x / y; const float u2 = u * u; return (u2 - 1.0f) / (u2 + 1.0f);
	}
	inline void vscopeMeters(VScopeFilters& f, const float* left, const float* right, std::size_t n, float envelopeCoeff, float stereoCoeff, float secondStereoFilterSpeed)
	{
		const float stereoPoles[2] = { stereoCoeff, static_cast<float>(std::pow(stereoCoeff, secondStereoFilterSpeed)) };
		const float mr = -0.70710678118654752440f, mi = 0.70710678118654752440f;
		for (std::size_t z = 0; z < n; ++z)
		{
			const float l = left[z], r = right[z];
			const float vX = l * mr - r * mi;
			const float vY = r * mi + l * mr;
			const float ph = cosTwiceAtan(vY, vX);
			const float lSquared = l * l, rSquared = r * r;
			f.envelope[0] = lSquared + envelopeCoeff * (f.envelope[0] - lSquared);
			f.envelope[1] = rSquared + envelopeCoeff * (f.envelope[1] - rSquared);
			for (int k = 0; k < 2; ++k)
			{
				f.balance[k][0] = lSquared + stereoPoles[k] * (f.balance[k][0] - lSquared);
				f.balance[k][1] = rSquared + stereoPoles[k] * (f.balance[k][1] - rSquared);
				f.phase[k] = ph + stereoPoles[k] * (f.phase[k] - ph);
			}
		}
	}

	// ---------------------------------------------------------------------------------------
	// Oscilloscope — channel views of SampleColourEvaluators.h:57-320: L, R, Mid = 0.5 (L + R),
	// Side = 0.5 (L - R)
	// ---------------------------------------------------------------------------------------
	enum class OscChannels : int { Left, Right, Mid, Side };
	inline float oscChannel(OscChannels m, float l, float r)
	{
		switch (m)
		{
		case OscChannels::Left: return l;
		case OscChannels::Right: return r;
		case OscChannels::Mid: return 0.5f * (l + r);
		default: return 0.5f * (l - r);
		}
	}

	// Lanczos branch of drawWavePlot — Oscilloscope/OscilloscopeRendering.cpp:790-891.
	// `x` is the evaluator's channel, n samples, newest last; samplePos as computed at :796-826.
	struct ResampleParams
	{
		double samplePos{ 0 };       // after :826
		double samplesPerPixel{ 1.0 / 16 };
		double left{ 0 }, right{ 1 }, inc{ 1.0 / 1024 };
	};

	inline std::size_t resampleCount(const ResampleParams& p)
	{
		std::size_t count = 0;
		double unitSpacePos = p.left;
		do { ++count; unitSpacePos += p.inc; } while (unitSpacePos < (p.right + p.inc));
		return count;
	}

	// out: count x (x, y) floats
	inline void lanczosResample(const ResampleParams& p, const float* x, std::ptrdiff_t n, float* out)
	{
		constexpr int KernelSize = 10, KernelBufferSize = KernelSize * 2 + 1;
		double samplePos = p.samplePos;
		double currentSample = std::floor(samplePos);
		// eval.startFrom(-floor(samplePos) - KernelSize): position relative to the end of the buffer
		std::ptrdiff_t cursor = n - static_cast<std::ptrdiff_t>(std::floor(samplePos)) - KernelSize;
		auto get = [&]() -> float { const float v = (cursor >= 0 && cursor < n) ? x[cursor] : 0.0f; ++cursor; return v; };
		float kernel[KernelBufferSize];
		for (auto& f : kernel) f = get();
		double unitSpacePos = p.left;
		std::size_t m = 0;
		do
		{
			double delta = currentSample - samplePos;
			while (delta > 1)
			{
				samplePos += 1; delta -= 1;
				std::rotate(kernel, kernel + 1, kernel + KernelBufferSize);
				kernel[KernelBufferSize - 1] = get();
			}
			const double v = dsp::lanczosFilterReal<double, float>(kernel, KernelBufferSize, KernelSize + delta, KernelSize);
			out[2 * m] = static_cast<float>(unitSpacePos);
			out[2 * m + 1] = static_cast<float>(v);
			++m;
			currentSample += p.samplesPerPixel;
			unitSpacePos += p.inc;
		} while (unitSpacePos < (p.right + p.inc));
	}

	// ZeroCrossingProcessor::process / PeakHoldProcessor::process — Oscilloscope/StreamPreprocessing.h:284-347
	struct TriggerState
	{
		double state{ 0 };
		bool isPeakHolding{ false };
		uint64_t crossOrigin{ 0 };
	};

	inline void zeroCrossingTrigger(TriggerState& s, const float* x, std::size_t n, double threshold, uint64_t steadyClock, std::vector<uint64_t>& peaks)
	{
		for (std::size_t count = 0; count < n; ++count)
		{
			const double sample = x[count];
			if (sample > 0 && s.state < 0) { s.isPeakHolding = true; s.crossOrigin = steadyClock + count; }
			if (s.isPeakHolding && sample > threshold) { s.isPeakHolding = false; peaks.push_back(s.crossOrigin); }
			s.state = sample;
		}
	}

	inline void peakHoldTrigger(TriggerState& s, const float* x, std::size_t n, double threshold, double hysteresis, uint64_t steadyClock, std::vector<uint64_t>& peaks)
	{
		for (std::size_t count = 0; count < n; ++count)
		{
			double sample = x[count];
			sample *= sample;
			const double delta = sample - s.state;
			if (delta < 0)
			{
				s.state *= 0.9999;
				s.state = std::max(threshold * threshold, s.state);
				if (s.isPeakHolding) { peaks.push_back(steadyClock + count - 1); s.isPeakHolding = false; }
			}
			else
			{
				if (delta > hysteresis * s.state) s.isPeakHolding = true;
				s.state = sample;
			}
		}
	}

	// TriggeringProcessor — Oscilloscope/StreamPreprocessing.h:40-226: turns the queue of trigger positions into display
	// windows.  `processIntoBackBuffer(samples)` (audio appended to the back buffer) and
	// `channelData.swapBuffers(cappedSize, -bufferedSamples)` (a window copied from the back to the front buffer) are
	// recorded as events instead of being executed; everything else is the reference's bookkeeping, with its types
	// (std::ceil / halfSize in double, the clocks in std::uint64_t, the mixed conversions of :146-191 as written).
	struct WindowEvent
	{
		enum Kind { Process = 0, Swap = 1 };
		int kind;
		uint64_t origin;   // Process: steadyClock before the samples; Swap: frontOrigin, the clock of the window's first sample
		uint64_t size;     // Process: samples; Swap: cappedSize
		uint64_t buffered; // bufferedSamples at the event (Swap: the offset handed to swapBuffers is -buffered)
	};

	struct TriggeringProcessor
	{
		double windowSize{ 0 };
		bool windowChanged{ false };
		uint64_t oldPeak{ 0 }, currentPeak{ 0 }, bufferedSamples{ 0 }, frontOrigin{ 0 }, steadyClock{ 0 };
		std::deque<uint64_t> peaks;
		bool isWorkingOnPeak{ false };

		// :45-52 (the part that concerns the window)
		void setWindowSize(double newWindowSize)
		{
			windowChanged = std::ceil(newWindowSize) != std::ceil(windowSize);
			windowSize = newWindowSize;
		}

		// :54-75
		void update(uint64_t currentSteadyClock)
		{
			steadyClock = currentSteadyClock;
			if (windowChanged)
			{
				windowChanged = false;
				if (isWorkingOnPeak && peaks.size()) peaks.pop_front(); // (the reference pops unconditionally)
				while (peaks.size() && peaks.front() < currentSteadyClock) peaks.pop_front();
				bufferedSamples = currentPeak = oldPeak = 0;
				frontOrigin = currentSteadyClock;
				isWorkingOnPeak = false;
			}
		}

		// :78-206
		void processMutating(std::size_t numSamples, std::vector<WindowEvent>& events)
		{
			if (frontOrigin + bufferedSamples < steadyClock)
			{
				frontOrigin = steadyClock;
				bufferedSamples = 0;
			}
			auto ceilingSize = std::ceil(windowSize);
			auto halfSize = ceilingSize / 2;
			auto processIntoBackBuffer = [&](auto samples)
			{
				events.push_back({ WindowEvent::Process, steadyClock, static_cast<uint64_t>(samples), bufferedSamples });
				numSamples -= samples;
				auto oldSamples = bufferedSamples;
				steadyClock += samples;
				bufferedSamples += samples;
				bufferedSamples = std::min<std::uint64_t>(bufferedSamples, ceilingSize + 1);
				frontOrigin += (oldSamples + samples) - bufferedSamples;
			};
			if (ceilingSize == 0 && peaks.size()) peaks.clear();
			while (numSamples != 0)
			{
				if (!peaks.size())
				{
					processIntoBackBuffer(numSamples);
					break;
				}
				else if (!isWorkingOnPeak)
				{
					isWorkingOnPeak = true;
					auto nextPeak = peaks.front();
					if (nextPeak >= steadyClock)
					{
						auto deltaToPeak = nextPeak - steadyClock;
						auto numSamplesToProcess = std::min<std::uint64_t>(numSamples, deltaToPeak + halfSize);
						processIntoBackBuffer(numSamplesToProcess);
						currentPeak = nextPeak;
					}
					else
					{
						currentPeak = nextPeak;
					}
				}
				std::uint64_t windowEnd;
				bool isPeakOutsideOfWindow = false;
				bool readyForBufferSwap = false;
				if (currentPeak - oldPeak < halfSize)
				{
					windowEnd = oldPeak + halfSize;
				}
				else
				{
					isPeakOutsideOfWindow = true;
					windowEnd = frontOrigin + bufferedSamples;
				}
				std::uint64_t peakWindowEnd = (isPeakOutsideOfWindow ? 1 : 0) + currentPeak + halfSize;
				std::uint64_t numSamplesToProcess = 0;
				auto missingBufferSamples = peakWindowEnd - std::min(peakWindowEnd, windowEnd);
				auto neededPreSamples = std::min<std::uint64_t>(halfSize, std::max<double>((currentPeak - oldPeak), halfSize) - halfSize);
				if (isPeakOutsideOfWindow)
				{
					numSamplesToProcess = std::min<std::uint64_t>(numSamples, missingBufferSamples);
					if (numSamplesToProcess > 0) processIntoBackBuffer(numSamplesToProcess);
					readyForBufferSwap = missingBufferSamples == numSamplesToProcess;
				}
				else
				{
					if (bufferedSamples >= missingBufferSamples)
					{
						readyForBufferSwap = true;
					}
					else
					{
						auto numRemaining = missingBufferSamples - bufferedSamples;
						numSamplesToProcess = std::min<std::uint64_t>(numSamples, numRemaining);
						if (numSamplesToProcess > 0) processIntoBackBuffer(numSamplesToProcess);
						readyForBufferSwap = numRemaining == numSamplesToProcess;
					}
				}
				if (readyForBufferSwap)
				{
					auto amount = (isPeakOutsideOfWindow ? halfSize : missingBufferSamples) + neededPreSamples;
					auto cappedSize = std::min<std::size_t>(bufferedSamples, std::ceil(amount + 1));
					events.push_back({ WindowEvent::Swap, frontOrigin, static_cast<uint64_t>(cappedSize), bufferedSamples });
					bufferedSamples -= std::min<std::uint64_t>(bufferedSamples, cappedSize);
					frontOrigin += cappedSize;
					oldPeak = currentPeak;
					isWorkingOnPeak = false;
					peaks.pop_front();
				}
			}
		}
	};

	// calculateFundamentalPeriod + calculateTriggeringOffset — Oscilloscope/OscilloscopeDSP.inl:61-308
	struct BinRecord
	{
		uint64_t index{ 0 };
		double value{ 0 }, offset{ 0 };
		double omega() const noexcept { return double(index) + offset; }
	};

	struct PeriodState
	{
		static constexpr std::size_t FilterSize = 8;
		std::array<BinRecord, FilterSize> median{};
		std::size_t medianPos{ 0 };
	};

	struct PeriodParams
	{
		double effectiveWindowSize{ 1024 }, sampleRate{ 48000 }, threshold{ 0 }, hysteresis{ 0 }, phaseOffsetDegrees{ 0 };
	};

	struct PeriodResult
	{
		BinRecord record;
		double fundamental{ 0 }, cycleSamples{ 0 }, phase{ 0 }, sampleOffset{ 0 };
	};

	inline PeriodResult spectralPeriod(PeriodState& st, const PeriodParams& p, const float* x, std::ptrdiff_t n)
	{
		constexpr std::size_t LookaheadSize = 8192, TransformSize = LookaheadSize;
		PeriodResult r;
		std::vector<std::complex<double>> transformBuffer(TransformSize), work(TransformSize);
		auto sampleAt = [&](std::ptrdiff_t i) -> double { return (i >= 0 && i < n) ? double(x[i]) : 0.0; };
		{
			// :88-95
			const auto offset = std::max<std::size_t>(static_cast<std::size_t>(std::ceil(p.effectiveWindowSize)), LookaheadSize);
			for (std::size_t i = 0; i < LookaheadSize; ++i) transformBuffer[i] = sampleAt(n - std::ptrdiff_t(offset) + std::ptrdiff_t(i));
		}
		dsp::UniFFT<double> fft(TransformSize); // signaldust::DustFFT_fwdDa: forward DFT, natural order (cpl decision)
		fft.forward(transformBuffer.data(), transformBuffer.data(), work.data());

		auto quadDelta = [&](std::size_t w) -> double
		{
			const auto x0 = transformBuffer[w], x1 = transformBuffer[w + 1], xm1 = transformBuffer[w == 0 ? 1 : w - 1];
			const auto denom = x0 * 2.0 - xm1 - x1;
			return (denom.real() + denom.imag()) != 0 ? std::real((xm1 - x1) / denom) : 0;
		};
		const double quarterSemitone = std::pow(2, 0.25 / 12.0) - 1;
		const double invHysteresis = 1 - p.hysteresis;
		BinRecord max{ 1, std::max(p.threshold * TransformSize / 6.0, std::abs(transformBuffer[1])), quadDelta(1) };
		for (std::size_t i = 2; i < (TransformSize >> 1); ++i)
		{
			BinRecord current{ i, std::abs(transformBuffer[i]), 0 };
			if (invHysteresis * current.value > max.value * 2)
			{
				if (max.omega() > 0)
				{
					current.offset = quadDelta(i);
					auto factor = current.omega() / max.omega();
					auto sensivity = current.value / max.value;
					if (invHysteresis * sensivity > 20) { max = current; continue; }
					if (std::abs(1 - factor) < invHysteresis * quarterSemitone) { max = current; continue; }
					auto multipleDeviation = std::abs(factor - std::floor(factor + 0.5));
					if (invHysteresis * std::abs(multipleDeviation) > quarterSemitone) max = current;
				}
				else
				{
					max = current;
					max.offset = quadDelta(max.index);
				}
			}
		}
		// :186-212 — median over the previous 8 picks.  std::nth_element leaves the choice among equal
		// indices to the implementation; restated as a stable sort (documented decision).
		auto localMedian = st.median;
		st.median[st.medianPos] = max;
		st.medianPos = (st.medianPos + 1) & (PeriodState::FilterSize - 1);
		std::stable_sort(localMedian.begin(), localMedian.end(), [](const BinRecord& a, const BinRecord& b) { return a.index < b.index; });
		const auto& oldMedianBin = localMedian[PeriodState::FilterSize >> 1];
		if (oldMedianBin.index != uint64_t(-1) && std::abs(max.omega() - oldMedianBin.omega()) > 0.5) max = oldMedianBin;
		r.record = max;
		double fundamental = p.sampleRate * max.omega() / TransformSize;
		r.fundamental = fundamental = std::max(5.0, fundamental);
		r.cycleSamples = p.sampleRate / fundamental;

		// calculateTriggeringOffset, Spectral mode (:249-306)
		const double tau = 6.283185307179586476925286766559;
		const double radians = tau * max.omega() / TransformSize;
		const double offsetReal = std::max<double>(double(LookaheadSize), p.effectiveWindowSize + r.cycleSamples);
		const auto offset = static_cast<std::size_t>(std::ceil(offsetReal));
		const double sampleDifference = double(offset) - (p.effectiveWindowSize + r.cycleSamples);
		std::vector<double> temporaryBuffer(LookaheadSize);
		for (std::size_t i = 0; i < LookaheadSize; ++i) temporaryBuffer[i] = sampleAt(n - std::ptrdiff_t(offset) + std::ptrdiff_t(i));
		auto z = dsp::goertzel(temporaryBuffer.data(), LookaheadSize, radians);
		const double rotation = -sampleDifference * radians;
		z *= std::complex<double>(std::cos(rotation), -std::sin(rotation));
		double phase = tau - std::arg(z);
		phase += max.offset * tau;
		phase -= 1.5707963267948966192313216916398;
		phase += tau * p.phaseOffsetDegrees / 360;
		phase = std::fmod(phase, tau);
		while (phase < 0) phase += tau;
		r.phase = phase;
		r.sampleOffset = (phase / tau) * p.sampleRate / r.fundamental - 1;
		return r;
	}
}
