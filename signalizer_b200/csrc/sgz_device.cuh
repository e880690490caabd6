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

// ------------------------------------------------------------------------------------------------
// Packed FP32 (Blackwell add/sub/mul/fma.f32x2 -> SASS FADD2 / FMUL2 / FFMA2).  The fused kernel
// carries two independent FFTs (the even- and odd-sample halves, E and O) through identical
// arithmetic, so it keeps them as register pairs: re = (E.re, O.re), im = (E.im, O.im).  Every
// butterfly add/sub and every twiddle multiply is then one packed instruction for both halves.
// ------------------------------------------------------------------------------------------------
namespace sgz
{
	typedef unsigned long long u64_t;
	__device__ __forceinline__ u64_t pk(float2 v) { return reinterpret_cast<const u64_t&>(v); }
	__device__ __forceinline__ float2 unpk(u64_t v) { return reinterpret_cast<const float2&>(v); }

	__device__ __forceinline__ float2 add2(float2 a, float2 b) { u64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return unpk(r); }
	__device__ __forceinline__ float2 sub2(float2 a, float2 b) { u64_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return unpk(r); }
	__device__ __forceinline__ float2 mul2(float2 a, float2 b) { u64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return unpk(r); }
	__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { u64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pk(a)), "l"(pk(b)), "l"(pk(c))); return unpk(r); }
	__device__ __forceinline__ float2 bc2(float s) { return make_float2(s, s); }

	// forward radix-4 on SoA pairs: (re, im)[0..3] -> X0..X3
	__device__ __forceinline__ void radix4p(float2& r0, float2& i0, float2& r1, float2& i1, float2& r2, float2& i2, float2& r3, float2& i3)
	{
		const float2 s02r = add2(r0, r2), s02i = add2(i0, i2), d02r = sub2(r0, r2), d02i = sub2(i0, i2);
		const float2 s13r = add2(r1, r3), s13i = add2(i1, i3), d13r = sub2(r1, r3), d13i = sub2(i1, i3);
		r0 = add2(s02r, s13r); i0 = add2(s02i, s13i);
		r2 = sub2(s02r, s13r); i2 = sub2(s02i, s13i);
		r1 = add2(d02r, d13i); i1 = sub2(d02i, d13r); // d02 - i d13
		r3 = sub2(d02r, d13i); i3 = add2(d02i, d13r); // d02 + i d13
	}

	// (re, im) *= (c - i s) with broadcast constants: re' = re c + im s, im' = im c - re s
	__device__ __forceinline__ void mulCS(float2& re, float2& im, float c, float s)
	{
		const float2 C = bc2(c), S = bc2(s);
		const float2 nr = fma2(im, S, mul2(re, C));
		const float2 ni = sub2(mul2(im, C), mul2(re, S));
		re = nr; im = ni;
	}

	template<int E> __device__ __forceinline__ void mulW16p(float2& re, float2& im)
	{
		constexpr float C1 = 0.92387953251128673848f, S1 = 0.38268343236508978178f, R = 0.70710678118654752440f;
		if constexpr (E == 0) {}
		else if constexpr (E == 1) mulCS(re, im, C1, S1);
		else if constexpr (E == 2) { const float2 a = add2(re, im), b = sub2(im, re); re = mul2(a, bc2(R)); im = mul2(b, bc2(R)); }
		else if constexpr (E == 3) mulCS(re, im, S1, C1);
		else if constexpr (E == 4) { const float2 t = re; re = im; im = sub2(bc2(0.f), t); }
		else if constexpr (E == 6) { const float2 a = sub2(im, re), b = add2(re, im); re = mul2(a, bc2(R)); im = mul2(b, bc2(-R)); }
		else if constexpr (E == 9) mulCS(re, im, -C1, -S1);
		else static_assert(E < 0, "unsupported W16 exponent");
	}

	__device__ __forceinline__ void dft16p(float2 (&re)[16], float2 (&im)[16])
	{
#pragma unroll
		for (int j = 0; j < 4; ++j) radix4p(re[j], im[j], re[j + 4], im[j + 4], re[j + 8], im[j + 8], re[j + 12], im[j + 12]);
		mulW16p<1>(re[5], im[5]); mulW16p<2>(re[9], im[9]); mulW16p<3>(re[13], im[13]);
		mulW16p<2>(re[6], im[6]); mulW16p<4>(re[10], im[10]); mulW16p<6>(re[14], im[14]);
		mulW16p<3>(re[7], im[7]); mulW16p<6>(re[11], im[11]); mulW16p<9>(re[15], im[15]);
#pragma unroll
		for (int m = 0; m < 4; ++m) radix4p(re[4 * m], im[4 * m], re[4 * m + 1], im[4 * m + 1], re[4 * m + 2], im[4 * m + 2], re[4 * m + 3], im[4 * m + 3]);
	}

	template<int O>
	__device__ __forceinline__ void dft8p(float2 (&re)[16], float2 (&im)[16])
	{
		constexpr float R = 0.70710678118654752440f;
		radix4p(re[O + 0], im[O + 0], re[O + 2], im[O + 2], re[O + 4], im[O + 4], re[O + 6], im[O + 6]);
		radix4p(re[O + 1], im[O + 1], re[O + 3], im[O + 3], re[O + 5], im[O + 5], re[O + 7], im[O + 7]);
		{
			const float2 a = add2(re[O + 3], im[O + 3]), b = sub2(im[O + 3], re[O + 3]);   // W8^1
			re[O + 3] = mul2(a, bc2(R)); im[O + 3] = mul2(b, bc2(R));
			const float2 t = re[O + 5]; re[O + 5] = im[O + 5]; im[O + 5] = sub2(bc2(0.f), t); // W8^2 = -i
			const float2 c = sub2(im[O + 7], re[O + 7]), d = add2(re[O + 7], im[O + 7]);   // W8^3
			re[O + 7] = mul2(c, bc2(R)); im[O + 7] = mul2(d, bc2(-R));
		}
#pragma unroll
		for (int m = 0; m < 4; ++m)
		{
			const float2 ar = re[O + 2 * m], ai = im[O + 2 * m], br = re[O + 2 * m + 1], bi = im[O + 2 * m + 1];
			re[O + 2 * m] = add2(ar, br); im[O + 2 * m] = add2(ai, bi);
			re[O + 2 * m + 1] = sub2(ar, br); im[O + 2 * m + 1] = sub2(ai, bi);
		}
	}

	// (re, im) *= w for a table twiddle stored pre-broadcast as float4 (wx, wx, wy, wy)
	__device__ __forceinline__ void mulTwiddleP(float2& re, float2& im, float4 w)
	{
		const float2 X = make_float2(w.x, w.y), Y = make_float2(w.z, w.w);
		const float2 nr = sub2(mul2(re, X), mul2(im, Y));
		const float2 ni = fma2(re, Y, mul2(im, X));
		re = nr; im = ni;
	}
}
