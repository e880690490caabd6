// Stand-in for <cpl/dsp/CComplexResonator.h> (see cpl/Common.h in this directory).  Also carries the other cpl::dsp
// names the path uses (WindowTypes, linearFilter, lanczosFilter, separateTransformsIPL); all forward to the oracle.
#pragma once
#include "../Common.h"
#include "../lib/uarray.h"

namespace cpl
{
	namespace dsp
	{
		using WindowTypes = sgzo::WindowTypes;

		template<typename C, typename V, typename X>
		inline C linearFilter(const V& v, X x)
		{
			typedef typename C::value_type S;
			return sgzo::dsp::linearFilter<S>(v.data(), static_cast<std::ptrdiff_t>(v.size()), static_cast<S>(x));
		}

		template<typename C, bool Precise, typename V, typename X>
		inline C lanczosFilter(const V& v, X x, int a)
		{
			typedef typename C::value_type S;
			return sgzo::dsp::lanczosFilter<S>(v.data(), static_cast<std::ptrdiff_t>(v.size()), static_cast<S>(x), a);
		}

		template<typename V>
		inline void separateTransformsIPL(V v)
		{
			// Separate / MidSide hand in csf.slice(0, N) (TransformDSP.inl:858), Phase the whole csf of N + 1 entries
			// (:646: the extra slot is filled by the caller afterwards).  cpl decision, as in the restated oracle: the
			// transform length is the even part of the view's length in both cases.
			sgzo::dsp::separateTransformsIPL(v.data(), v.size() & ~std::size_t(1));
		}

		template<typename T, std::size_t Channels>
		class CComplexResonator
		{
		public:
			typedef sgzo::ResonatorConstant<T> Constant;

			void resetState() { impl.resetState(); }

			template<typename V>
			void resonateReal(const Constant& c, const uarray<T>& work, std::size_t channels, std::size_t numSamples)
			{
				const T* chans[1] = { work.data() };
				impl.resonateReal(c, chans, channels, numSamples);
			}
			template<typename V, std::size_t N>
			void resonateReal(const Constant& c, const std::array<uarray<T>, N>& work, std::size_t channels, std::size_t numSamples)
			{
				const T* chans[N];
				for (std::size_t i = 0; i < N; ++i) chans[i] = work[i].data();
				impl.resonateReal(c, chans, channels, numSamples);
			}
			template<typename ISA>
			void getWholeWindowedState(const Constant& c, WindowTypes window, uarray<T> out, std::size_t outChannels, std::size_t numFilters)
			{
				impl.getWholeWindowedState(c, window, out.data(), outChannels, numFilters);
			}

			sgzo::ComplexResonator<T> impl;
		};
	}
}
