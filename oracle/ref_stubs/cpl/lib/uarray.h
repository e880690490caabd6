// Stand-in for <cpl/lib/uarray.h> (see cpl/Common.h in this directory): a non-owning view with slice / reinterpret.
#pragma once
#include "../Common.h"

namespace cpl
{
	template<typename T>
	class uarray
	{
	public:
		typedef T value_type;
		uarray() = default;
		uarray(T* data, std::size_t size) : ptr(data), count(size) {}
		// view of const from view of non-const
		template<typename U, typename = std::enable_if_t<std::is_same_v<const U, T>>>
		uarray(const uarray<U>& other) : ptr(other.data()), count(other.size()) {}

		T& operator[](std::size_t i) const { return ptr[i]; }
		T* data() const { return ptr; }
		T* begin() const { return ptr; }
		T* end() const { return ptr + count; }
		std::size_t size() const { return count; }
		uarray slice(std::size_t offset, std::size_t n) const { return uarray(ptr + offset, n); }
		template<typename Y>
		uarray<Y> reinterpret() const { return uarray<Y>(reinterpret_cast<Y*>(ptr), count * sizeof(T) / sizeof(Y)); }

	private:
		T* ptr{ nullptr };
		std::size_t count{ 0 };
	};

	template<typename T, typename A>
	inline uarray<T> as_uarray(std::vector<T, A>& v) { return uarray<T>(v.data(), v.size()); }
	template<typename T, typename A>
	inline uarray<const T> as_uarray(const std::vector<T, A>& v) { return uarray<const T>(v.data(), v.size()); }
}
