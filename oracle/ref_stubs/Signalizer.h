// Stand-in for Source/Signalizer.h so that the reference's Source/Spectrum/{TransformDSP.inl, TransformPair.h,
// TransformConstant.h} compile unmodified (oracle/build_ref.sh copies those three next to this file, into
// oracle/_ref/src/Spectrum/).  TEST INFRASTRUCTURE ONLY.  Declares just what those files name:
//   AudioStream::{DataType, AudioBufferView, bufferIndices}  (cpl::AudioStream 2-segment ring views, TransformDSP.inl:65-88)
//   SpectrumChannels, UComplexFilter                         (Source/Common/CommonSignalizer.h:495-591: values / layout)
//   FloatColour, ColourRotation, ParameterSet::ParameterView, juce::ColourGradient   (named by TransformConstant.h, off the path)
//   cpl::ParameterWindowDesignValue                          (generateWindow<T> -> the oracle's window decision)
#pragma once

#include "cpl/Common.h"
#include "cpl/lib/uarray.h"
#include "cpl/dsp/CComplexResonator.h"

namespace juce
{
	struct ColourGradient
	{
		template<typename C> void addColour(double, const C&) {}
	};
}

namespace cpl
{
	// the window designer parameter: holds the choice, generateWindow<T>(kernel, W) returns windowKernelScale
	template<typename View>
	struct ParameterWindowDesignValue
	{
		sgzo::WindowTypes type{ sgzo::WindowTypes::Hann };
		sgzo::WindowShape shape{ sgzo::WindowShape::Periodic };
		template<typename T, typename Vector>
		T generateWindow(Vector& kernel, std::size_t W) { return sgzo::dsp::generateWindow<T>(kernel.data(), W, type, shape); }
	};
}

namespace Signalizer
{
	struct AudioStream
	{
		typedef float DataType;
		static constexpr std::size_t bufferIndices = 2;
		typedef sgzo::AudioBufferView AudioBufferView;
	};

	struct ParameterSet { struct ParameterView {}; };
	struct FloatColour { float r{ 0 }, g{ 0 }, b{ 0 }, a{ 0 }; };
	typedef std::array<FloatColour, 4> ColourRotation;

	// CommonSignalizer.h:495-539 (the numeric values are what the C ABI's SGZ_CH_* constants mirror)
	enum class SpectrumChannels { Left, Right, Merge, Mid = Merge, Side, OffsetForMono = Side, Phase, Separate, MidSide, Complex, End };

	// CommonSignalizer.h:541-591: three views of one pair of scalars
	template<typename Scalar>
	union UComplexFilter
	{
		UComplexFilter() : real(0), imag(0) {}
		UComplexFilter(const std::complex<Scalar>& c) : real(c.real()), imag(c.imag()) {}
		UComplexFilter& operator=(const std::complex<Scalar>& c) noexcept { real = c.real(); imag = c.imag(); return *this; }
		struct { Scalar real, imag; };
		struct { Scalar magnitude, phase; };
		struct { Scalar leftMagnitude, rightMagnitude; };
		operator std::complex<Scalar>() const noexcept { return std::complex<Scalar>(real, imag); }
	};
}
