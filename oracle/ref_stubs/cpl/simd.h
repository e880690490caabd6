// Stand-in for <cpl/simd.h> (see cpl/Common.h in this directory): only consts<T> is named on the path.
#pragma once
#include "Common.h"

namespace cpl
{
	namespace simd
	{
		template<typename T>
		struct consts
		{
			static constexpr T half = T(0.5);
			static constexpr T one = T(1);
			static constexpr T zero = T(0);
		};
	}
}
