// Stand-in for Source/Spectrum/SpectrumParameters.h (the real one is the GUI parameter set): only the enums and the
// constant that TransformDSP.inl / TransformPair.h / TransformConstant.h name (SpectrumParameters.h:48-78).
// TEST INFRASTRUCTURE ONLY (oracle/build_ref.sh).
#pragma once

#include "../Signalizer.h"

namespace Signalizer
{
	struct SpectrumContent
	{
		enum LineGraphs { None = -2, Transform = -1, LineMain = 0, LineSecond, LineEnd };
		enum class BinInterpolation { None, Linear, Lanczos };
		enum class DisplayMode { LineGraph, ColourSpectrum };
		enum class TransformAlgorithm { FFT, RSNT };
		enum class ViewScaling { Linear, Logarithmic };
		static const std::size_t numSpectrumColours = 5;
	};
}
