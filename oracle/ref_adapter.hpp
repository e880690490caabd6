// ref_adapter.hpp — drives the REFERENCE's own Spectrum transform code (Source/Spectrum/TransformDSP.inl,
// TransformPair.h, TransformConstant.h, copied at build time into oracle/_ref/src/Spectrum/ by oracle/build_ref.sh and
// compiled against the stand-in headers under oracle/ref_stubs/) through the same C entry points as the restated oracle.
//
// TEST INFRASTRUCTURE ONLY.  The classes here only rename / re-shape arguments so that sgz_oracle_capi.cpp can hold
// either implementation; every statement of prepareTransform, doTransform, mapToLinearSpace, postProcessStdTransform,
// audioEntryPoint, resonatingDispatch, remapFrequencies, generateSlopeMap and setStorage that runs is the reference's.
// The cpl primitives underneath (FFT, filters, window, resonator) are the oracle's "cpl decisions" (cpl is absent).
#pragma once

#include <algorithm>
#include <array>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <optional>
#include <vector>

// the harness reads and writes the peak-hold states, which the reference keeps private to TransformPair
#define private public
#include "Spectrum/TransformDSP.inl"
#undef private

namespace sgzref
{
	using Signalizer::SpectrumChannels;
	typedef Signalizer::SpectrumContent::BinInterpolation BinInterpolation;
	typedef Signalizer::SpectrumContent::DisplayMode DisplayMode;
	typedef Signalizer::SpectrumContent::TransformAlgorithm TransformAlgorithm;
	typedef Signalizer::SpectrumContent::ViewScaling ViewScaling;
	using sgzo::WindowTypes;
	using sgzo::WindowShape;
	using sgzo::AudioBufferView;
	using sgzo::AudioPair;
	namespace dsp { using sgzo::dsp::windowCoefficientCount; }
	namespace Math { using sgzo::Math::confineTo; }

	// the reference's ISA tag: only ISA::V (the SIMD vector type handed to the resonator) is named on the path
	struct ScalarISA { typedef float V; };

	template<typename T>
	class TransformConstant : public Signalizer::TransformConstant<T>
	{
		typedef Signalizer::TransformConstant<T> Base;
	public:
		void remapFrequencies(double left, double right, ViewScaling scaling, double minFreq)
		{
			cpl::Utility::Bounds<double> view; view.left = left; view.right = right;
			Base::remapFrequencies(view, scaling, minFreq);
		}
		void generateSlopeMap(double a, double b) { Base::generateSlopeMap(cpl::PowerSlopeValue::PowerFunction{ a, b }); }
		void regenerateWindowKernel(WindowTypes type, WindowShape shape)
		{
			cpl::ParameterWindowDesignValue<Signalizer::ParameterSet::ParameterView> designer;
			designer.type = type; designer.shape = shape;
			Base::regenerateWindowKernel(designer);
			this->dspWindow = type; // Spectrum.cpp:583-588 sets both from the same parameter
		}
	};

	template<typename T>
	class TransformPair : public Signalizer::TransformPair<T>
	{
		typedef Signalizer::TransformPair<T> Base;
		typedef TransformConstant<T> Constant;
	public:
		bool prepareTransform(const Constant& c, const AudioPair& views) { return Base::prepareTransform(c, views); }
		bool prepareTransform(const Constant& c, const AudioPair& views, std::array<const float*, 2> prelim, std::size_t n)
		{
			return Base::prepareTransform(c, views, { const_cast<float*>(prelim[0]), const_cast<float*>(prelim[1]) }, n);
		}
		void mapToLinearSpace(const Constant& c) { Base::template mapToLinearSpace<ScalarISA>(c); }
		const std::complex<T>* getRawFFT(const Constant& c) { return Base::getRawFFT(c).data(); }
		const T* getTransformResult(const Constant& c, std::size_t* n)
		{
			auto r = Base::getTransformResult(c);
			*n = r.size();
			return r.data();
		}
		void resonatingDispatch(const Constant& c, std::array<const float*, 2> buffer, std::size_t n)
		{
			Base::template resonatingDispatch<ScalarISA>(c, { const_cast<float*>(buffer[0]), const_cast<float*>(buffer[1]) }, n);
		}
		void audioEntryPoint(const Constant& c, const AudioPair* views, std::array<const float*, 2> buffer, std::size_t n)
		{
			std::optional<AudioPair> v;
			if (views) v = *views;
			Base::template audioEntryPoint<ScalarISA>(c, v, { const_cast<float*>(buffer[0]), const_cast<float*>(buffer[1]) }, n);
		}
	};

	// first / second scalar of a result or state entry (UComplexFilter: real / imag = magnitude / phase = left / right)
	template<typename S> inline S first(const Signalizer::UComplexFilter<S>& v) { return v.real; }
	template<typename S> inline S second(const Signalizer::UComplexFilter<S>& v) { return v.imag; }
	template<typename S> inline void assign(Signalizer::UComplexFilter<S>& v, S a, S b) { v.real = a; v.imag = b; }
	template<typename T> inline sgzo::ComplexResonator<T>& resonatorOf(TransformPair<T>& p) { return p.cresonator.impl; }
}
