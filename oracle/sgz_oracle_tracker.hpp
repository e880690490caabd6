// sgz_oracle_tracker.hpp — TEST INFRASTRUCTURE ONLY (see oracle/sgz_oracle.hpp): CPU restatement of the
// peak search of Spectrum::drawFrequencyTracking (Source/Spectrum/SpectrumRendering.cpp:301-469, 494-519)
// and of SmoothedPeakState (Source/Spectrum/Spectrum.h:405-459).  PARITY UNPINNED (no reference tests or
// fixtures exist; cpl is absent) — the cpl pieces the reference calls here are restated as decisions:
//   * cpl::Math::round<T>(x)            -> (T)std::floor(x + 0.5)
//   * cpl::Math::square(complex)         -> re*re + im*im in float
//   * cpl::Math::UnityScale::linear(x, lo, hi) -> x * (hi - lo) + lo; Inv::linear -> (y - lo) / (hi - lo)
//   * frequencyGraph.fractionToCoordTransformed(fraction of Nyquist) -> the inverse of
//     TransformConstant::remapFrequencies (TransformConstant.h:125-182) for non-Complex configurations
//   * cpl::dsp::SmoothedParameterState<double, 8>: eight cascaded one-pole smoothers,
//     design(ms, rate) = exp(-1 / (ms * 0.001 * rate)), process(pole, x): s[i] = in + pole * (s[i] - in)
#pragma once
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstdint>

#include "../include/sgz.h"

namespace sgz_oracle
{
	struct TrackerView
	{
		// what the tracker reads from Spectrum::state / Constant
		int graph;               // SGZ_TRACK_*
		bool fft;                // state.algo == FFT
		int configuration;       // SGZ_CH_*
		int binPolation;         // SGZ_BIN_*
		int viewScale;           // 0 linear, 1 logarithmic (as sgz_spectrum_config.view_scale)
		double viewLeft, viewRight, minLogFreq;
		double sampleRate;       // config->sampleRate
		std::size_t transformSize, windowSize, axisPoints;
		double windowScale;      // constant.windowKernelScale
		double lowDbs, highDbs;  // getDBs()
		const float* mappedFrequencies; // [axisPoints]
		const float* slopeMap;          // [axisPoints]
	};

	template<typename T> inline T roundTo(double x) { return static_cast<T>(std::floor(x + 0.5)); }
	template<typename T> inline T confine(T x, T lo, T hi) { return x < lo ? lo : (x > hi ? hi : x); }

	inline double fractionToCoord(const TrackerView& v, double fraction)
	{
		const double viewSize = v.viewRight - v.viewLeft;
		double arg;
		if (v.viewScale == 0) arg = fraction;
		else
		{
			const double end = v.sampleRate / 2;
			arg = std::log(fraction * end / v.minLogFreq) / std::log(end / v.minLogFreq);
		}
		return (arg - v.viewLeft) / viewSize * double(v.axisPoints - 1);
	}

	inline double smoothDesign(double ms, double rate) { return std::exp(-1.0 / (ms * 0.001 * rate)); }
	inline void smoothProcess(double (&s)[8], double pole, double in)
	{
		for (double& x : s) { x = in + pole * (x - in); in = x; }
	}

	// results: [axisPoints][2] floats (UComplex, .leftMagnitude first); rawFFT: [transformSize (+1)][2] floats
	inline void trackPeak(const TrackerView& v, const float* results, const float* rawFFT, double mouseFraction, double smoothingMs, double fps,
		sgz_tracker_state* st, sgz_tracker_result* out)
	{
		double peakFrequency = 0, peakDBs = 0, peakFraction = 0, peakX = 0, peakDeviance = 0, peakSlope = 0;
		const double sampleRate = v.sampleRate;
		const double nearbyFractionToConsider = 0.03, precisionError = 0.001, interpolationError = 0.01;
		mouseFraction = confine(mouseFraction, 0.0, 1.0);
		int graphN = v.graph;
		std::ptrdiff_t peakOffset = 0;
		const bool rawPath = !(v.configuration == SGZ_CH_COMPLEX || !(v.fft && graphN == SGZ_TRACK_TRANSFORM));
		if (!rawPath)
		{
			// SpectrumRendering.cpp:306-374
			if (graphN == SGZ_TRACK_TRANSFORM) graphN = SGZ_TRACK_LINE_MAIN; // (the caller passes that graph's results)
			const std::size_t N = v.axisPoints;
			const auto pivot = roundTo<std::size_t>(N * mouseFraction);
			const auto range = roundTo<std::size_t>(N * nearbyFractionToConsider);
			const std::size_t lowerBound = range > pivot ? 0 : pivot - range;
			const std::size_t higherBound = range + pivot > N ? N : range + pivot;
			auto mag = [&](std::size_t i) { return results[2 * i]; };
			std::size_t peak = lowerBound;
			for (std::size_t i = lowerBound; i < higherBound; ++i) if (mag(peak) < mag(i)) peak = i; // std::max_element: first maximum
			if (peak == lowerBound && lowerBound != 0)
			{
				while (true)
				{
					const std::size_t next = peak - 1;
					if (next == 0) break;
					else if (mag(next) < mag(peak)) break;
					else peak = next;
				}
			}
			else if (peak == higherBound - 1)
			{
				while (true)
				{
					const std::size_t next = peak + 1;
					if (next == N) break;
					else if (mag(next) < mag(peak)) break;
					else peak = next;
				}
			}
			peakOffset = std::ptrdiff_t(peak);
			peakFrequency = v.mappedFrequencies[peak];
			const bool offsetIsEnd = peak == v.axisPoints - 1;
			// float - float, widened on assignment
			peakDeviance = v.mappedFrequencies[offsetIsEnd ? peak : peak + 1] - v.mappedFrequencies[offsetIsEnd ? peak - 1 : peak];
			if (v.fft && v.binPolation != SGZ_BIN_LANCZOS) peakDeviance = std::max(peakDeviance, 0.5 * v.transformSize / N);
			peakX = double(peakOffset);
			peakSlope = v.slopeMap[peak];
			const double peakFractionY = results[2 * peak];
			peakDBs = peakFractionY * (v.highDbs - v.lowDbs) + v.lowDbs;
		}
		else
		{
			// SpectrumRendering.cpp:376-469: search the original FFT
			const std::ptrdiff_t N = std::ptrdiff_t(v.transformSize), points = std::ptrdiff_t(v.axisPoints);
			auto lowerBound = roundTo<std::ptrdiff_t>(points * (mouseFraction - nearbyFractionToConsider));
			lowerBound = roundTo<std::ptrdiff_t>((N * v.mappedFrequencies[confine<std::ptrdiff_t>(lowerBound, 0, points - 1)] / sampleRate));
			auto higherBound = roundTo<std::ptrdiff_t>(points * (mouseFraction + nearbyFractionToConsider));
			higherBound = roundTo<std::ptrdiff_t>((N * v.mappedFrequencies[confine<std::ptrdiff_t>(higherBound, 0, points - 1)] / sampleRate));
			lowerBound = confine<std::ptrdiff_t>(lowerBound, 0, N);
			higherBound = confine<std::ptrdiff_t>(higherBound, 0, N);
			// the reference indexes source[N] when higherBound == N (one past getRawFFT's N elements, the extra
			// element audioMemory carries, TransformPair.h:164-167); the buffer passed here holds N + 1 entries
			auto sq = [&](std::ptrdiff_t i) { const float re = rawFFT[2 * i], im = rawFFT[2 * i + 1]; return re * re + im * im; };
			auto cabs = [&](std::ptrdiff_t i, float scale) { return std::abs(std::complex<float>(rawFFT[2 * i], rawFFT[2 * i + 1]) * scale); };
			std::ptrdiff_t peak = lowerBound;
			for (std::ptrdiff_t i = lowerBound; i < higherBound + 1; ++i) if (sq(peak) < sq(i)) peak = i;
			if (peak == lowerBound && lowerBound != 0)
			{
				while (true)
				{
					const std::ptrdiff_t next = peak - 1;
					if (next == 0) break;
					else if (sq(next) < sq(peak)) break;
					else peak = next;
				}
			}
			else if (peak == higherBound - 1)
			{
				while (true)
				{
					const std::ptrdiff_t next = peak + 1;
					if (next == N) break; // source.end()
					else if (sq(next) < sq(peak)) break;
					else peak = next;
				}
			}
			peakOffset = peak;
			const float invSize = static_cast<float>(v.windowScale / (v.windowSize * 0.5));
			// 20 * log10(float) stays float (int * float)
			const float alpha = 20 * std::log10(cabs(peakOffset == 0 ? 0 : peakOffset - 1, invSize));
			const float beta = 20 * std::log10(cabs(peakOffset, invSize));
			const float gamma = 20 * std::log10(cabs(peakOffset == N ? peakOffset : peakOffset + 1, invSize));
			const double phi = 0.5 * (alpha - gamma) / (alpha - 2 * beta + gamma);
			peakFraction = 2 * (peakOffset + (std::isnormal(phi) ? phi : 0)) / double(N);
			peakFrequency = 0.5 * peakFraction * sampleRate;
			peakDBs = beta - 0.25 * (alpha - gamma) * phi;
			if (!std::isnormal(peakDBs))
				peakDBs = 20 * std::log10(std::abs(std::complex<float>(rawFFT[2 * peakOffset], rawFFT[2 * peakOffset + 1])) / (N * 0.5));
			peakX = fractionToCoord(v, peakFraction);
			peakSlope = v.slopeMap[confine<std::ptrdiff_t>(roundTo<std::ptrdiff_t>(peakX), 0, points - 1)];
			peakDBs += 20 * std::log10(peakSlope);
			const double normalizedDeviation = (1.0 - peakFraction) / N;
			peakDeviance = 2 * interpolationError * normalizedDeviation * sampleRate + precisionError;
		}
		// :494-519
		const bool peakIsComplex = peakFrequency > sampleRate * 0.5;
		if (peakIsComplex) peakFrequency = sampleRate - peakFrequency;
		if (peakDBs > 1000) peakDBs = INFINITY;
		else if (peakDBs < -1000) peakDBs = -INFINITY;
		out->peak_offset = peakOffset;
		out->searched_raw_fft = rawPath ? 1 : 0;
		out->peak_is_complex = peakIsComplex ? 1 : 0;
		out->peak_frequency = peakFrequency; out->peak_dbs = peakDBs; out->peak_deviance = peakDeviance;
		out->peak_slope = peakSlope; out->peak_x = peakX; out->peak_fraction = peakFraction;
		out->smoothed_frequency = peakFrequency; out->smoothed_dbs = peakDBs;
		if (st)
		{
			// SmoothedPeakState::clearNonNormals / design / process (Spectrum.h:410-444)
			if (!std::isnormal(st->current_peak)) st->current_peak = 0;
			if (!std::isnormal(st->frequency_state[7])) for (double& x : st->frequency_state) x = 0;
			if (!std::isnormal(st->peak_dbs_state[7])) for (double& x : st->peak_dbs_state) x = 0;
			const double ms = smoothingMs;
			const double peakPole = smoothDesign(ms * 10, fps), filterPole = smoothDesign(ms / 5, fps);
			const double peak = std::pow(10, peakDBs / 20);
			st->current_peak *= peakPole;
			if (peak > st->current_peak)
			{
				st->current_peak = 1.2 * (peak / peakPole);
				st->current_frequency = peakFrequency;
				st->current_peak_dbs = peakDBs;
			}
			smoothProcess(st->frequency_state, filterPole, st->current_frequency);
			smoothProcess(st->peak_dbs_state, filterPole, st->current_peak_dbs);
			out->smoothed_frequency = st->frequency_state[7];
			out->smoothed_dbs = st->peak_dbs_state[7];
		}
	}
}
