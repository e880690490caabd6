// Stand-in for <cpl/Common.h> so that the reference's own Source/Spectrum/{TransformDSP.inl, TransformPair.h,
// TransformConstant.h} compile unmodified (oracle/build_ref.sh).  TEST INFRASTRUCTURE ONLY.
//
// The reference's numeric library cpl (bitbucket.org/Mayae/cpl @ fa8e21e1) is absent from /root/reference; every cpl
// primitive below forwards to the "cpl decision" of the same name in oracle/sgz_oracle.hpp, so oracle/_ref executes
// the reference's in-tree control flow and arithmetic on top of exactly the primitives the restated oracle uses.
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
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "sgz_oracle.hpp"

#define CPL_RUNTIME_ASSERTION(expr) do { if (!(expr)) throw std::runtime_error("CPL_RUNTIME_ASSERTION: " #expr); } while (0)
#define CPL_RUNTIME_EXCEPTION(msg) throw std::runtime_error(msg)

namespace cpl
{
	typedef std::ptrdiff_t ssize_t;

	namespace Types
	{
		// cpl decision (same as the restated oracle): the "fast" integers are plain int
		typedef int fint_t;
		typedef int fsint_t;
		typedef unsigned fuint_t;
	}

	// cpl::aligned_vector<T, alignment>: a std::vector with an aligned allocator; alignment does not change results
	template<typename T, std::size_t Alignment>
	using aligned_vector = std::vector<T>;

	namespace Math
	{
		inline std::size_t nextPow2Inc(std::size_t x) { return sgzo::Math::nextPow2Inc(x); }
		template<typename T, typename A, typename B> inline T confineTo(T v, A lo, B hi) { return sgzo::Math::confineTo<T>(v, static_cast<T>(lo), static_cast<T>(hi)); }
		template<typename T> inline T square(const std::complex<T>& z) { return sgzo::Math::square(z); }
		template<typename T> inline T square(T x) { return x * x; }
		template<typename T> inline T dbToFraction(T db) { return static_cast<T>(sgzo::Math::dbToFraction(static_cast<double>(db))); }
	}

	// cpl::CPeakFilter<T>: only `.pole` is read on the path
	template<typename T>
	using CPeakFilter = sgzo::CPeakFilter<T>;

	namespace Utility
	{
		template<typename T>
		struct Bounds
		{
			T left{}, right{}, top{}, bottom{};
			T dist() const { return right - left; }
		};
	}

	struct PowerSlopeValue
	{
		struct PowerFunction { double a, b; };
	};
}
