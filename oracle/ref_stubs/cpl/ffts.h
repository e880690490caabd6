// Stand-in for <cpl/ffts.h> (see cpl/Common.h in this directory): UniFFT forwards to the oracle's FFT.
#pragma once
#include "Common.h"
#include "lib/uarray.h"

namespace cpl
{
	namespace dsp
	{
		template<typename C> class UniFFT;

		template<typename S>
		class UniFFT<std::complex<S>>
		{
		public:
			UniFFT() = default;
			UniFFT(std::size_t n) : impl(n) {}
			void forward(uarray<std::complex<S>> in, uarray<std::complex<S>> out, uarray<std::complex<S>> work) const
			{
				impl.forward(in.data(), out.data(), work.data());
			}
		private:
			sgzo::dsp::UniFFT<S> impl;
		};
	}
}
