// sgz_oracle_colour.hpp — TEST INFRASTRUCTURE ONLY (see oracle/sgz_oracle.hpp): CPU restatement of the
// oscilloscope's per-sample DSP, Oscilloscope::StreamState::audioProcessing
// (Source/Oscilloscope/OscilloscopeDSP.inl:426-710: RMS envelope, 3-band crossover, colour smoothing,
// accumulateColour) and Oscilloscope::runPeakFilter (:712-886).  PARITY UNPINNED: the reference has no
// tests for it and the crossover itself lives in the absent cpl library.  cpl pieces restated as decisions:
//   * cpl::dsp::LinkwitzRileyNetwork<float, 3>: 4th-order Linkwitz-Riley tree.  Each crossover is a pair of
//     2nd-order Butterworth sections (bilinear transform, Q = 1/sqrt 2, designed in double, stored as float)
//     applied twice; x -> LP1^2 -> allpass(f2) = LP2^2 + HP2^2 -> band 0;  x -> HP1^2 -> LP2^2 -> band 1,
//     -> HP2^2 -> band 2.  Sections are transposed direct form II, evaluated without fused multiply-add.
//   * cpl::dsp::SmoothedParameterState<float, 1>::design(ms, rate) = exp(-1 / (ms * 0.001 * rate))
//   * UPixel::lerp(key, t): per component (including alpha) c + (key - c) * t in float, truncated
//   * static_cast<uint8_t>(NaN) (all bands silent: 0 * inf) -> 0
//   * consts<V>::sign_mask in runPeakFilter clears the sign bit (|x|); V holds 8 floats (AVX)
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>

#include "../include/sgz.h"

namespace sgz_oracle
{
	struct Section { float b0, b1, b2, a1, a2; };
	struct CrossoverCoeffs { Section lp[2], hp[2]; }; // [crossover]

	inline void designButterworth(double fNorm, Section& lp, Section& hp)
	{
		const double w = 2.0 * 3.14159265358979323846 * fNorm, cs = std::cos(w), sn = std::sin(w);
		const double alpha = sn / (2.0 * 0.70710678118654752440), a0 = 1.0 + alpha;
		lp = Section{ float((1 - cs) * 0.5 / a0), float((1 - cs) / a0), float((1 - cs) * 0.5 / a0), float(-2 * cs / a0), float((1 - alpha) / a0) };
		hp = Section{ float((1 + cs) * 0.5 / a0), float(-(1 + cs) / a0), float((1 + cs) * 0.5 / a0), float(-2 * cs / a0), float((1 - alpha) / a0) };
	}

	// ChannelData::tuneCrossOver(300, 3000, sampleRate): design({ (float)(300 / fs), (float)(3000 / fs) })
	inline CrossoverCoeffs designCrossover(double sampleRate)
	{
		CrossoverCoeffs c;
		designButterworth(static_cast<float>(300.0 / sampleRate), c.lp[0], c.hp[0]);
		designButterworth(static_cast<float>(3000.0 / sampleRate), c.lp[1], c.hp[1]);
		return c;
	}

	inline float runSection(const Section& s, float* z, float x)
	{
		const float y = s.b0 * x + z[0];
		z[0] = (s.b1 * x - s.a1 * y) + z[1];
		z[1] = s.b2 * x - s.a2 * y;
		return y;
	}

	// network state: 12 sections x 2: [LP1 LP1 | HP1 HP1 | apLP2 apLP2 | apHP2 apHP2 | LP2 LP2 | HP2 HP2]
	inline void runNetwork(const CrossoverCoeffs& c, float* z, float x, float (&bands)[3])
	{
		const float low = runSection(c.lp[0], z + 2, runSection(c.lp[0], z + 0, x));
		const float rest = runSection(c.hp[0], z + 6, runSection(c.hp[0], z + 4, x));
		const float apl = runSection(c.lp[1], z + 10, runSection(c.lp[1], z + 8, low));
		const float aph = runSection(c.hp[1], z + 14, runSection(c.hp[1], z + 12, low));
		bands[0] = apl + aph;
		bands[1] = runSection(c.lp[1], z + 18, runSection(c.lp[1], z + 16, rest));
		bands[2] = runSection(c.hp[1], z + 22, runSection(c.hp[1], z + 20, rest));
	}

	inline uint8_t toByte(float v) { return v != v ? uint8_t(0) : static_cast<uint8_t>(v); }

	// accumulateColour (:474-497)
	inline void accumulateColour(const float (&state)[3], const float (&colours)[3][3], const uint8_t (&key)[4], float blend, uint8_t* out)
	{
		float red = 0, green = 0, blue = 0;
		for (int i = 0; i < 3; ++i)
		{
			red += state[i] * colours[i][0];
			green += state[i] * colours[i][1];
			blue += state[i] * colours[i][2];
		}
		const float invMax = 255.0f / std::max(red, std::max(blue, green));
		const uint8_t px[4] = { toByte(red * invMax), toByte(green * invMax), toByte(blue * invMax), 255 };
		for (int k = 0; k < 4; ++k) out[k] = toByte(float(px[k]) + (float(key[k]) - float(px[k])) * blend);
	}

	inline void smoothBands(const float (&bands)[3], float (&states)[3], float pole)
	{
		for (int i = 0; i < 3; ++i)
		{
			const float input = bands[i] * bands[i];
			states[i] = input + pole * (states[i] - input);
		}
	}

	// one channel pair, numSamples samples; st[0] = left/mid state, st[1] = right/side state
	// rgba: 4 planes [n][4] in the order left, right, mid, side.  Returns envelopeGain (or 0 when the mode is off).
	inline double audioProcessing(const sgz_scope_colour_config& cfg, const float* left, const float* right, std::size_t n, sgz_scope_colour_state* st, uint8_t* rgba)
	{
		const CrossoverCoeffs coeffs = designCrossover(cfg.sample_rate);
		const float pole = static_cast<float>(std::exp(-1.0 / (cfg.colour_smoothing_ms * 0.001 * cfg.sample_rate)));
		const float envelopeCoeff = static_cast<float>(std::exp(-1.0 / (cfg.envelope_window * cfg.sample_rate)));
		const float blend = 1 - static_cast<float>(cfg.colouring_blend);
		float filterEnv[2] = { st[0].envelope, st[1].envelope };
		if (cfg.envelope_mode != SGZ_ENVELOPE_NONE)
		{
			auto pole1 = [&](float& e, float sample) { e = sample + envelopeCoeff * (e - sample); };
			switch (cfg.channel_mode)
			{
			case SGZ_OSC_RIGHT: for (std::size_t i = 0; i < n; ++i) pole1(filterEnv[0], right[i] * right[i]); filterEnv[1] = filterEnv[0]; break;
			case SGZ_OSC_LEFT: for (std::size_t i = 0; i < n; ++i) pole1(filterEnv[0], left[i] * left[i]); filterEnv[1] = filterEnv[0]; break;
			case SGZ_OSC_MID:
				for (std::size_t i = 0; i < n; ++i) { const float m = 0.5f * (left[i] + right[i]); pole1(filterEnv[0], m * m); }
				filterEnv[1] = filterEnv[0]; break;
			case SGZ_OSC_SIDE:
				for (std::size_t i = 0; i < n; ++i) { const float s = 0.5f * (left[i] - right[i]); pole1(filterEnv[0], s * s); }
				filterEnv[1] = filterEnv[0]; break;
			case SGZ_OSC_SEPARATE:
				for (std::size_t i = 0; i < n; ++i) { pole1(filterEnv[0], left[i] * left[i]); pole1(filterEnv[1], right[i] * right[i]); }
				break;
			case SGZ_OSC_MIDSIDE:
				for (std::size_t i = 0; i < n; ++i)
				{
					const float l = left[i], r = right[i];
					const float mid = 0.5f * ((l + r) * (l + r)), side = 0.5f * ((l - r) * (l - r));
					pole1(filterEnv[0], mid); pole1(filterEnv[1], side);
				}
				break;
			}
		}
		uint8_t* planes[4] = { rgba, rgba + 4 * n, rgba + 8 * n, rgba + 12 * n };
		for (std::size_t i = 0; i < n; ++i)
		{
			float lb[3], rb[3], mb[3], sb[3];
			runNetwork(coeffs, st[0].network, left[i], lb);
			runNetwork(coeffs, st[1].network, right[i], rb);
			smoothBands(lb, st[0].smooth_filters, pole);
			smoothBands(rb, st[1].smooth_filters, pole);
			for (int k = 0; k < 3; ++k) { mb[k] = lb[k] + rb[k]; sb[k] = lb[k] - rb[k]; }
			smoothBands(mb, st[0].aux_smooth_filter, pole);
			smoothBands(sb, st[1].aux_smooth_filter, pole);
			accumulateColour(st[0].smooth_filters, cfg.colours, cfg.key_left, blend, planes[0] + 4 * i);
			accumulateColour(st[1].smooth_filters, cfg.colours, cfg.key_right, blend, planes[1] + 4 * i);
			accumulateColour(st[0].aux_smooth_filter, cfg.colours, cfg.key_left, blend, planes[2] + 4 * i);
			accumulateColour(st[1].aux_smooth_filter, cfg.colours, cfg.key_right, blend, planes[3] + 4 * i);
		}
		double gain = 0;
		if (cfg.envelope_mode == SGZ_ENVELOPE_RMS)
		{
			float start = std::sqrt(filterEnv[0]);
			for (int c = 0; c < 2; ++c) start = std::max(start, std::sqrt(filterEnv[c]));
			gain = 1.0 / start;
			st[0].envelope = filterEnv[0];
			st[1].envelope = filterEnv[1];
		}
		return gain;
	}

	// runPeakFilter (:712-886) for one stereo pair; env[2] = smoothEnvelopeState(0 / 1).  Returns state.autoGain.
	inline double runPeakFilter(int channelMode, const float* left, const float* right, std::size_t n, double coeff, float* env)
	{
		const std::size_t stop = n - (n & 7);
		auto fold = [&](auto value) { float m = 0; for (std::size_t i = 0; i < stop; ++i) m = std::max(std::fabs(value(i)), m); return m; };
		if (channelMode <= SGZ_OSC_SIDE)
		{
			float hi = 0;
			switch (channelMode)
			{
			case SGZ_OSC_LEFT: hi = fold([&](std::size_t i) { return left[i]; }); break;
			case SGZ_OSC_RIGHT: hi = fold([&](std::size_t i) { return right[i]; }); break;
			case SGZ_OSC_MID: hi = fold([&](std::size_t i) { return (left[i] + right[i]) * 0.5f; }); break;
			default: hi = fold([&](std::size_t i) { return (left[i] - right[i]) * 0.5f; }); break;
			}
			const double h = hi;
			env[0] = static_cast<float>(std::max(env[0] * coeff, h * h));
			env[1] = static_cast<float>(std::max(env[1] * coeff, h * h));
		}
		else if (channelMode == SGZ_OSC_SEPARATE)
		{
			// the running maximum is not reset between channels (the reference keeps one vLMax)
			const float hl = fold([&](std::size_t i) { return left[i]; });
			env[0] = std::max<float>(env[0] * coeff, hl * hl);
			const float hr = std::max(hl, fold([&](std::size_t i) { return right[i]; }));
			env[1] = std::max<float>(env[1] * coeff, hr * hr);
		}
		else
		{
			const double hm = fold([&](std::size_t i) { return (left[i] + right[i]) * 0.5f; });
			const double hs = fold([&](std::size_t i) { return (left[i] - right[i]) * 0.5f; });
			env[0] = static_cast<float>(std::max(env[0] * coeff, hm * hm));
			env[1] = static_cast<float>(std::max(env[1] * coeff, hs * hs));
		}
		float start = std::sqrt(env[0]);
		for (int c = 0; c < 2; ++c) start = std::max(start, std::sqrt(env[c]));
		return 1.0 / start;
	}
	// VectorScope::runPeakFilter (Source/Vectorscope/VectorscopeRendering.cpp:826-889) for one stereo pair.
	// env[2] = processor->filters.envelope (relaxed_atomic<float>), gain = processor->envelopeGain (in/out).
	inline void runVectorscopePeakFilter(const float* left, const float* right, std::size_t numSamples, float envelopeCoeff, double openGLDeltaTime,
		float* env, double* gain)
	{
		double currentEnvelope = 1;
		{
			// :840-842
			double power = numSamples * openGLDeltaTime;
			double coeff = std::pow(envelopeCoeff, power);
			// :845-868: vLMax / vRMax over whole vectors of 8 floats, |x| through the sign mask
			float lmax[8] = { 0, 0, 0, 0, 0, 0, 0, 0 }, rmax[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
			const std::size_t loopIncrement = 8;
			std::size_t stop = numSamples - (numSamples & (loopIncrement - 1));
			for (std::size_t i = 0; i < stop; i += loopIncrement)
				for (std::size_t k = 0; k < loopIncrement; ++k)
				{
					lmax[k] = std::max(std::fabs(left[i + k]), lmax[k]);
					rmax[k] = std::max(std::fabs(right[i + k]), rmax[k]);
				}
			// :874-880
			double highestLeft = *std::max_element(lmax, lmax + 8);
			double highestRight = *std::max_element(rmax, rmax + 8);
			env[0] = static_cast<float>(std::max(env[0] * coeff, highestLeft * highestLeft));
			env[1] = static_cast<float>(std::max(env[1] * coeff, highestRight * highestRight));
			// :882: std::sqrt of the float states
			currentEnvelope = 1.0 / std::max(std::sqrt(env[0]), std::sqrt(env[1]));
		}
		// :885-888
		if (std::isnormal(currentEnvelope)) *gain = currentEnvelope;
	}
}
