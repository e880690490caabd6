// sgz_device.cuh — device-side building blocks for the sm_100a spectrum kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace sgz
{
	__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
	__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
	__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
	__device__ __forceinline__ float2 cscale(float2 a, float s) { return make_float2(a.x * s, a.y * s); }
	__device__ __forceinline__ float cnorm(float2 a) { return a.x * a.x + a.y * a.y; }
	__device__ __forceinline__ float cabs(float2 a) { return sqrtf(a.x * a.x + a.y * a.y); }

	// forward radix-4 butterfly, in place: (a0,a1,a2,a3) -> (X0,X1,X2,X3)
	__device__ __forceinline__ void radix4(float2& a0, float2& a1, float2& a2, float2& a3)
	{
		const float2 s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), d13 = csub(a1, a3);
		a0 = cadd(s02, s13);
		a2 = csub(s02, s13);
		a1 = make_float2(d02.x + d13.y, d02.y - d13.x); // d02 - i d13
		a3 = make_float2(d02.x - d13.y, d02.y + d13.x); // d02 + i d13
	}

	// multiply by W16^e = exp(-2 pi i e / 16) for the exponents a 4x4 radix-16 needs
	template<int E> __device__ __forceinline__ float2 mulW16(float2 a)
	{
		constexpr float C1 = 0.92387953251128673848f; // cos(pi/8)
		constexpr float S1 = 0.38268343236508978178f; // sin(pi/8)
		constexpr float R = 0.70710678118654752440f;
		if constexpr (E == 0) return a;
		else if constexpr (E == 1) return make_float2(a.x * C1 + a.y * S1, a.y * C1 - a.x * S1);
		else if constexpr (E == 2) return make_float2((a.x + a.y) * R, (a.y - a.x) * R);
		else if constexpr (E == 3) return make_float2(a.x * S1 + a.y * C1, a.y * S1 - a.x * C1);
		else if constexpr (E == 4) return make_float2(a.y, -a.x);
		else if constexpr (E == 6) return make_float2((a.y - a.x) * R, -(a.x + a.y) * R);
		else if constexpr (E == 9) return make_float2(-a.x * C1 - a.y * S1, a.x * S1 - a.y * C1);
		else { static_assert(E < 0, "unsupported W16 exponent"); return a; }
	}

	// 16-point forward DFT in place; output X[k] ends up in v[perm16(k)]
	__host__ __device__ constexpr int perm16(int k) { return 4 * (k & 3) + (k >> 2); }

	__device__ __forceinline__ void dft16(float2 (&v)[16])
	{
#pragma unroll
		for (int j = 0; j < 4; ++j) radix4(v[j], v[j + 4], v[j + 8], v[j + 12]);
		// t[j][m] sits at v[j + 4m]; twiddle by W16^(j m)
		v[1 + 4] = mulW16<1>(v[1 + 4]); v[1 + 8] = mulW16<2>(v[1 + 8]); v[1 + 12] = mulW16<3>(v[1 + 12]);
		v[2 + 4] = mulW16<2>(v[2 + 4]); v[2 + 8] = mulW16<4>(v[2 + 8]); v[2 + 12] = mulW16<6>(v[2 + 12]);
		v[3 + 4] = mulW16<3>(v[3 + 4]); v[3 + 8] = mulW16<6>(v[3 + 8]); v[3 + 12] = mulW16<9>(v[3 + 12]);
#pragma unroll
		for (int m = 0; m < 4; ++m) radix4(v[4 * m], v[4 * m + 1], v[4 * m + 2], v[4 * m + 3]);
	}

	// 8-point forward DFT in place on v[o..o+8); output X[k] ends up in v[o + perm8(k)]
	__host__ __device__ constexpr int perm8(int k) { return 2 * (k & 3) + (k >> 2); }

	template<int O>
	__device__ __forceinline__ void dft8(float2 (&v)[16])
	{
		constexpr float R = 0.70710678118654752440f;
		// n = j + 2 i: radix-4 over i for j = 0, 1; t[j][m] at v[j + 2m]
		radix4(v[O + 0], v[O + 2], v[O + 4], v[O + 6]);
		radix4(v[O + 1], v[O + 3], v[O + 5], v[O + 7]);
		// twiddle t[1][m] by W8^m
		{
			float2 a = v[O + 3]; v[O + 3] = make_float2((a.x + a.y) * R, (a.y - a.x) * R);
			a = v[O + 5]; v[O + 5] = make_float2(a.y, -a.x);
			a = v[O + 7]; v[O + 7] = make_float2((a.y - a.x) * R, -(a.x + a.y) * R);
		}
		// radix-2 over j: X[m] = t0 + t1 (at v[2m]), X[m+4] = t0 - t1 (at v[2m+1])
#pragma unroll
		for (int m = 0; m < 4; ++m)
		{
			const float2 a = v[O + 2 * m], b = v[O + 2 * m + 1];
			v[O + 2 * m] = cadd(a, b);
			v[O + 2 * m + 1] = csub(a, b);
		}
	}

	// streaming 128-bit global load that does not allocate in L1 (audio is read once per frame)
	__device__ __forceinline__ float4 ldg_stream_f4(const float4* p)
	{
		float4 r;
		asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
		return r;
	}
	// table loads through the read-only path that the compiler may not hoist out of the frame loop
	// (hoisting 16 window + 46 twiddle values per thread would blow the 128-register budget)
	__device__ __forceinline__ float2 ldg_f2(const float2* p)
	{
		float2 r;
		asm volatile("ld.global.nc.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
		return r;
	}
	__device__ __forceinline__ void stg_stream_f2(float2* p, float2 v)
	{
		asm volatile("st.global.cs.v2.f32 [%0], {%1, %2};" :: "l"(p), "f"(v.x), "f"(v.y) : "memory");
	}
}
