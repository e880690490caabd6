// sgz_tables.hpp — host-side derivation of everything TransformConstant<T> holds
// (reference: Source/Spectrum/TransformConstant.h:81-180, Spectrum.cpp:365-394,489,571-595)
// plus the integer/weight tables that turn the sequential loops of mapToLinearSpace
// (TransformDSP.inl:562-639, :666-850, :875-984, :995-1096) into table-driven parallel work.
//
// Product code: this is NOT the oracle and shares no source with it; parity tests compare the
// two bit for bit on the integer tables.
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include "../../include/sgz.h"

namespace sgz
{
	constexpr int kLanczosSize = 5;               // lanczosFilterSize, TransformDSP.inl:514
	constexpr int kLanczosTaps = 2 * kLanczosSize; // floor(x)-a+1 .. floor(x)+a
	constexpr int kLanczosStride = 12;             // weights per point in the tables (10 taps + 2 zeros, float4 aligned)

	struct Tables
	{
		// sizes
		uint32_t N{ 0 }, W{ 0 }, P{ 0 }, numBins{ 0 };
		uint32_t stateChannels{ 1 };
		float sampleRate{ 0 }, topFrequency{ 0 }, freqToBin{ 0 };
		float windowKernelScale{ 0 }, invSize{ 0 };
		// postProcess scalars (TransformDSP.inl:1308-1314)
		float deltaYRecip{ 0 }, minFracRecip{ 0 }, lowerClip{ 0 };
		float pole[2]{ 0, 0 }, phasePole[2]{ 0, 0 };
		uint32_t breakPoint{ 0 };

		std::vector<float> window;        // [N], zero beyond W
		std::vector<float> mappedFreq;    // [P]
		std::vector<float> slope;         // [P]

		// per display point (all [P])
		std::vector<int32_t> mode;        // 0 interpolate, 1 max-of-bins
		std::vector<int32_t> first;       // interpolate: nearest (None) / floor(pos); max: first bin scanned
		std::vector<int32_t> count;       // max: bins scanned
		std::vector<int32_t> fallback;    // max: initial maxBin
		std::vector<float> posL, posR;    // fractional csf positions: f*freqToBin and N - that (float32)
		std::vector<int32_t> baseL, baseR; // Lanczos: first tap index floor(pos) - a + 1
		std::vector<float> wL, wR;        // Lanczos weights [P][kLanczosStride]
		// Phase configuration: number of low csf indices already replaced by their magnitude when
		// the magnitude pass evaluates point x (TransformDSP.inl:709-727, :759-776), and after it
		std::vector<int32_t> phaseNormPos; // [P]
		int32_t phaseNormFinal{ 0 };

		// smallest L such that every interpolation tap index i of every point satisfies
		// i < L or i > N - L (fused kernel keeps those entries as magnitudes, the rest squared)
		uint32_t linLimit{ 0 };
		// interpolated display points x >= safeFrom have every (Lanczos) tap of both channels inside [0, N]
		uint32_t safeFrom{ 0 };
		// true when the fused kernel reproduces the staged result for this configuration
		bool fusedOk{ false };
		// cosine-sum window spanning the whole transform in its periodic shape: window[n] = sum_k winPoly[k] cos^k(2 pi n / N)
		bool winPolyOk{ false };
		float winPoly[5]{ 0, 0, 0, 0, 0 };

		uint32_t resonatorVectors{ 1 };
		std::vector<float> resCoeffReal, resCoeffImag; // [V][P]
		std::vector<float> resGain;                    // [P]
		std::vector<float> resWindowTaps;              // [V] c_m
	};

	// returns SGZ_OK or SGZ_EINVAL
	int buildTables(const sgz_spectrum_config& cfg, Tables& out);

	std::size_t nextPow2Inc(std::size_t x);
	double lanczosKernel(double d, int a);
}
