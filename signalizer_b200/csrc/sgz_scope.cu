// sgz_scope.cu — sm_100a kernels and C ABI of the secondary paths: Vectorscope (polar / Lissajous
// vertices, point-density histogram, stereo meters) and Oscilloscope (Lanczos resampler, stream
// triggers, spectral period search).  References are relative to the reference checkout.
//
// Bit-exactness: histogram counts and trigger positions must equal the CPU oracle's exactly, so the
// float/double arithmetic that decides them is written with explicit round-to-nearest intrinsics
// (__fmul_rn, __fadd_rn, ...) that the compiler may not contract into FMAs.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/sgz.h"
#include "sgz_device.cuh"

namespace sgz
{
	int fail(int code, const char* fmt, ...); // sgz_spectrum.cu
}

#define SGZS_CUDA(expr)                                                                                         \
	do {                                                                                                        \
		cudaError_t e__ = (expr);                                                                               \
		if (e__ != cudaSuccess)                                                                                 \
			return sgz::fail(SGZ_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
	} while (0)

struct sgz_context
{
	int device{ 0 };
	cudaStream_t stream{ nullptr };
	bool ownStream{ false };
	void* stage[4]{ nullptr, nullptr, nullptr, nullptr };
	size_t stageBytes[4]{ 0, 0, 0, 0 };
	// x positions of the resampler's vertices for the last (left, right, inc): the reference accumulates
	// unitSpacePos += inc serially (OscilloscopeRendering.cpp:863-886), which left + m * inc only
	// reproduces for dyadic inc, so the host replays the accumulation once per view and caches it
	float* xTable{ nullptr };
	size_t xCapacity{ 0 };
	uint64_t xCount{ 0 };
	double xKey[3]{ 0, 0, 0 };
};

namespace sgz
{
	struct CtxGuard
	{
		int prev{ -1 };
		bool ok{ true };
		explicit CtxGuard(int dev)
		{
			if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
			if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
		}
		~CtxGuard() { if (prev >= 0) cudaSetDevice(prev); }
	};

	static int ensure(sgz_context* c, int slot, size_t need, void** out)
	{
		if (c->stageBytes[slot] < need)
		{
			if (c->stage[slot]) { cudaFree(c->stage[slot]); c->stage[slot] = nullptr; c->stageBytes[slot] = 0; }
			SGZS_CUDA(cudaMalloc(&c->stage[slot], std::max<size_t>(need, 256)));
			c->stageBytes[slot] = std::max<size_t>(need, 256);
		}
		*out = c->stage[slot];
		return SGZ_OK;
	}

	static int validateAudio(const float* audio, int layout, uint32_t nStreams, uint64_t n, uint64_t stride)
	{
		if (!audio) return fail(SGZ_EINVAL, "null audio");
		if (layout != SGZ_LAYOUT_PLANAR && layout != SGZ_LAYOUT_INTERLEAVED) return fail(SGZ_EINVAL, "bad layout");
		if (nStreams == 0) return fail(SGZ_EINVAL, "zero streams");
		if (nStreams > 1 && stride < 2 * n) return fail(SGZ_EINVAL, "stream_stride smaller than one stream");
		if (layout == SGZ_LAYOUT_INTERLEAVED && ((stride & 1) || (reinterpret_cast<uintptr_t>(audio) & 7)))
			return fail(SGZ_EINVAL, "interleaved audio must be 8-byte aligned with an even stream_stride");
		return SGZ_OK;
	}

	static int stageIn(sgz_context* c, const float* audio, int mem, uint32_t nStreams, uint64_t n, uint64_t stride, const float** dptr)
	{
		if (mem == SGZ_MEM_DEVICE) { *dptr = audio; return SGZ_OK; }
		const size_t floats = size_t(stride) * (nStreams - 1) + size_t(n) * 2;
		void* p = nullptr;
		const int rc = ensure(c, 0, floats * sizeof(float), &p);
		if (rc != SGZ_OK) return rc;
		SGZS_CUDA(cudaMemcpyAsync(p, audio, floats * sizeof(float), cudaMemcpyHostToDevice, c->stream));
		*dptr = static_cast<const float*>(p);
		return SGZ_OK;
	}

	__device__ __forceinline__ void loadLR(const float* base, int layout, uint64_t n, uint64_t i, float& l, float& r)
	{
		if (layout == SGZ_LAYOUT_INTERLEAVED) { const float2 v = reinterpret_cast<const float2*>(base)[i]; l = v.x; r = v.y; }
		else { l = base[i]; r = base[n + i]; }
	}

	__device__ __forceinline__ float oscChannel(int mode, float l, float r)
	{
		switch (mode)
		{
		case SGZ_OSC_LEFT: return l;
		case SGZ_OSC_RIGHT: return r;
		case SGZ_OSC_MID: return __fmul_rn(0.5f, __fadd_rn(l, r));
		default: return __fmul_rn(0.5f, __fsub_rn(l, r));
		}
	}

	// ------------------------------------------------------------------------------------------
	// Vectorscope
	// ------------------------------------------------------------------------------------------
	// (sin, cos)(atan(X / Y)) as (t, 1) / sqrt(1 + t^2), correctly rounded operation by operation
	// (cpl::simd::atan / sincos are unknown SSE polynomials; DESIGN.md "cpl decisions")
	__device__ __forceinline__ void sincosOfAtanRatio(float X, float Y, float& sx, float& cx)
	{
		const float aX = fabsf(X), aY = fabsf(Y);
		if (aX == 0.0f && aY == 0.0f) { sx = 0.0f; cx = 1.0f; return; }
		if (aY >= aX)
		{
			const float t = __fdiv_rn(X, Y);
			const float h = __fsqrt_rn(__fadd_rn(1.0f, __fmul_rn(t, t)));
			cx = __fdiv_rn(1.0f, h);
			sx = __fdiv_rn(t, h);
		}
		else
		{
			const float u = __fdiv_rn(Y, X);
			const float h = __fsqrt_rn(__fadd_rn(1.0f, __fmul_rn(u, u)));
			cx = __fdiv_rn(fabsf(u), h);
			const float m = __fdiv_rn(1.0f, h);
			sx = (signbit(X) != signbit(Y)) ? -m : m;
		}
	}

	__device__ __forceinline__ float3 vscopeVertex(const sgz_vscope_config& p, float left, float right, uint64_t i, uint64_t numSamples)
	{
		float3 v;
		if (p.polar)
		{
			const float cosineRotation = -0.70710678118654752440f, sineRotation = 0.70710678118654752440f;
			const float length = fmaxf(fabsf(left), fabsf(right));
			const float vY = __fsub_rn(__fmul_rn(left, cosineRotation), __fmul_rn(right, sineRotation));
			const float vX = __fadd_rn(__fmul_rn(left, sineRotation), __fmul_rn(right, cosineRotation));
			float sx, cx;
			if (left == 0.0f && right == 0.0f) { sx = 0.0f; cx = 1.0f; }
			else sincosOfAtanRatio(vX, vY, sx, cx);
			const float fadePerSample = __fdiv_rn(1.0f, float(numSamples));
			v = make_float3(__fmul_rn(sx, length), __fmul_rn(cx, length), __fsub_rn(__fmul_rn(float(i), fadePerSample), 1.0f));
		}
		else
		{
			const long long denom = (long long)numSamples - 1;
			const float sampleFade = __fdiv_rn(1.0f, float(denom < 1 ? 1 : int(denom)));
			v = make_float3(right, left, __fsub_rn(__fmul_rn(float(i), sampleFade), 1.0f));
		}
		v.x = __fmul_rn(v.x, p.gain); v.y = __fmul_rn(v.y, p.gain);
		if (p.polar && p.scale_polar) v.y = __fsub_rn(__fmul_rn(v.y, 2.0f), 1.0f);
		return v;
	}

	__global__ void k_vscope_vertices(sgz_vscope_config p, const float* __restrict__ audio, int layout, uint64_t n, uint64_t stride, float* __restrict__ xyz)
	{
		const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
		const uint32_t s = blockIdx.y;
		if (i >= n) return;
		float l, r;
		loadLR(audio + uint64_t(s) * stride, layout, n, i, l, r);
		const float3 v = vscopeVertex(p, l, r, i, n);
		float* o = xyz + (uint64_t(s) * n + i) * 3;
		o[0] = v.x; o[1] = v.y; o[2] = v.z;
	}

	// one CTA column per stream slice; counts go straight to the L2-resident histogram with
	// no-return reductions (RED.ADD), the natural fit for a 1 MiB-per-stream target
	__global__ void k_vscope_histogram(sgz_vscope_config p, const float* __restrict__ audio, int layout, uint64_t n, uint64_t stride, uint32_t* __restrict__ hist)
	{
		const uint32_t s = blockIdx.y;
		const float hw = __fmul_rn(0.5f, float(p.width)), hh = __fmul_rn(0.5f, float(p.height));
		uint32_t* h = hist + uint64_t(s) * p.width * p.height;
		for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x)
		{
			float l, r;
			loadLR(audio + uint64_t(s) * stride, layout, n, i, l, r);
			const float3 v = vscopeVertex(p, l, r, i, n);
			const float fx = floorf(__fmul_rn(__fadd_rn(v.x, 1.0f), hw)), fy = floorf(__fmul_rn(__fadd_rn(v.y, 1.0f), hh));
			if (!(fx >= 0.0f && fx < float(p.width) && fy >= 0.0f && fy < float(p.height))) continue;
			atomicAdd(h + size_t(fy) * p.width + size_t(fx), 1u);
		}
	}

	__device__ __forceinline__ float cosTwiceAtan(float y, float x)
	{
		if (x == 0.0f && y == 0.0f) return 0.0f;
		if (fabsf(x) >= fabsf(y)) { const float t = __fdiv_rn(y, x); const float t2 = __fmul_rn(t, t); return __fdiv_rn(__fsub_rn(1.0f, t2), __fadd_rn(1.0f, t2)); }
		const float u = __fdiv_rn(x, y); const float u2 = __fmul_rn(u, u); return __fdiv_rn(__fsub_rn(u2, 1.0f), __fadd_rn(u2, 1.0f));
	}

	// one warp per stream: lanes compute the per-sample inputs of 32 samples in parallel, lanes 0..7
	// then each run one of the eight one-pole recurrences over them in order (bit-exact with the
	// sequential reference; the filters are y <- x + p (y - x))
	__global__ void k_vscope_meters(const float* __restrict__ audio, int layout, uint64_t n, uint64_t stride, uint32_t nStreams,
		float envelopeCoeff, float pole0, float pole1, float* __restrict__ filters /* [S][8] */)
	{
		const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
		if (warp >= nStreams) return;
		const float* base = audio + uint64_t(warp) * stride;
		// filter slots: 0,1 envelope L/R; 2,3 balance slow L/R; 4,5 balance fast L/R; 6 phase slow; 7 phase fast
		float y = lane < 8 ? filters[warp * 8 + lane] : 0.f;
		const float pole = lane < 2 ? envelopeCoeff : ((lane == 2 || lane == 3 || lane == 6) ? pole0 : pole1);
		const int src = lane < 6 ? (lane & 1) : 2; // which input the slot consumes: lSquared, rSquared, phase
		const float mr = -0.70710678118654752440f, mi = 0.70710678118654752440f;
		for (uint64_t n0 = 0; n0 < n; n0 += 32)
		{
			float l = 0.f, r = 0.f;
			if (n0 + lane < n) loadLR(base, layout, n, n0 + lane, l, r);
			const float lSq = __fmul_rn(l, l), rSq = __fmul_rn(r, r);
			const float vX = __fsub_rn(__fmul_rn(l, mr), __fmul_rn(r, mi));
			const float vY = __fadd_rn(__fmul_rn(r, mi), __fmul_rn(l, mr));
			const float ph = cosTwiceAtan(vY, vX);
			const int count = int(min(uint64_t(32), n - n0));
			for (int j = 0; j < count; ++j)
			{
				const float a = __shfl_sync(0xffffffffu, lSq, j), b = __shfl_sync(0xffffffffu, rSq, j), c = __shfl_sync(0xffffffffu, ph, j);
				const float x = src == 0 ? a : (src == 1 ? b : c);
				y = __fadd_rn(x, __fmul_rn(pole, __fsub_rn(y, x)));
			}
		}
		if (lane < 8) filters[warp * 8 + lane] = y;
	}

	// ------------------------------------------------------------------------------------------
	// Oscilloscope: Lanczos resampler (a = 10, 21-tap window)
	// ------------------------------------------------------------------------------------------
	struct ResampleArgs
	{
		const float* audio; int layout; uint64_t n; uint64_t stride; int channel;
		const double* samplePos; // [S] device
		double spp; uint64_t count;
		const float* xs; // [count] device, float(unitSpacePos) per vertex
		float* out;
	};

	// cos(pi k / 10), sin(pi k / 10), k = 0..10
	__constant__ double kCos10[11] = { 1.0, 0.95105651629515357212, 0.80901699437494742410, 0.58778525229247312917, 0.30901699437494742410, 0.0,
		-0.30901699437494742410, -0.58778525229247312917, -0.80901699437494742410, -0.95105651629515357212, -1.0 };
	__constant__ double kSin10[11] = { 0.0, 0.30901699437494742410, 0.58778525229247312917, 0.80901699437494742410, 0.95105651629515357212, 1.0,
		0.95105651629515357212, 0.80901699437494742410, 0.58778525229247312917, 0.30901699437494742410, 0.0 };

	// One thread per output vertex.  The 20 Lanczos weights sinc(d) sinc(d / 10), d = frac + j, share their
	// transcendentals: sin(pi d) = (-1)^j sin(pi frac) and sin(pi d / 10) follows from sincospi(frac / 10)
	// by the angle-addition formula with the constants above (two transcendental calls per vertex
	// instead of forty).
	__global__ void k_scope_resample(ResampleArgs a)
	{
		const uint64_t m = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
		const uint32_t s = blockIdx.y;
		if (m >= a.count) return;
		constexpr int KernelSize = 10, KernelBufferSize = 21;
		const double samplePos0 = a.samplePos[s];
		const double fl0 = floor(samplePos0);
		// closed form of the sequential loop at OscilloscopeRendering.cpp:863-886: after m steps
		// currentSample = floor(samplePos0) + m spp, and the window has advanced k whole samples so
		// that delta = currentSample - (samplePos0 + k) <= 1
		const double currentSample = fl0 + double(m) * a.spp;
		const double d0 = currentSample - samplePos0;
		const double k = d0 > 1.0 ? ceil(d0 - 1.0) : 0.0;
		const double delta = currentSample - (samplePos0 + k);
		const long long base = (long long)a.n - (long long)fl0 - KernelSize + (long long)k; // index of kernel[0]
		const double x = double(KernelSize) + delta;
		const double fx = floor(x);
		const int ib = int(fx);
		const double frac = x - fx;
		const float* src = a.audio + uint64_t(s) * a.stride;
		const double S1 = sinpi(frac);
		double sa, ca;
		sincospi(frac * 0.1, &sa, &ca);
		const double invPi2x10 = 10.0 / (3.1415926535897932384626433832795 * 3.1415926535897932384626433832795);
		double acc = 0.0;
#pragma unroll
		for (int j = 9; j >= -10; --j) // tap i = ib - j ascending, distance d = frac + j
		{
			const int i = ib - j;
			if (i < 0 || i >= KernelBufferSize) continue;
			const long long idx = base + i;
			float v = 0.f;
			if (idx >= 0 && idx < (long long)a.n)
			{
				float l, r;
				loadLR(src, a.layout, a.n, uint64_t(idx), l, r);
				v = oscChannel(a.channel, l, r);
			}
			const double d = frac + double(j);
			double w;
			if (d == 0.0) w = 1.0;
			else
			{
				const int aj = j < 0 ? -j : j;
				const double cj = kCos10[aj], sj = j < 0 ? -kSin10[aj] : kSin10[aj];
				const double s10 = sa * cj + ca * sj;          // sin(pi d / 10)
				const double s1 = (aj & 1) ? -S1 : S1;         // sin(pi d)
				w = s1 * s10 * invPi2x10 / (d * d);
			}
			acc += double(v) * w;
		}
		float* o = a.out + (uint64_t(s) * a.count + m) * 2;
		o[0] = a.xs[m];
		o[1] = float(acc);
	}

	// Polyphase variant for integer oversampling ratios R = 1 / spp (BASELINE config 4: R = 16): in steady
	// state the fractional position of output m depends only on m mod R (and on the stream's
	// samplePos), so a CTA evaluates the R x 20 double weights once into shared memory and every
	// vertex is 20 double FMAs.  A vertex whose delta differs from its phase's (the first R outputs,
	// before the window starts sliding) takes the direct evaluation.
	constexpr int kMaxPhases = 64;
#ifndef SGZ_POLY_VPT
#define SGZ_POLY_VPT 8
#endif
#ifndef SGZ_POLY_MINB
#define SGZ_POLY_MINB 2
#endif
#ifndef SGZ_POLY_GROUPS
#define SGZ_POLY_GROUPS 8
#endif
	constexpr int kPolyVerticesPerThread = SGZ_POLY_VPT;       // vertices a thread keeps in flight at once
	constexpr int kPolyGroups = SGZ_POLY_GROUPS;               // such sweeps per CTA: the block prologue is paid once for all of them
	constexpr int kPolyBlockVertices = 256 * kPolyVerticesPerThread * kPolyGroups;
	__device__ __forceinline__ double lanczosWeight10(double frac, int ib, int t)
	{
		// weight of tap i = ib - 9 + t at distance d = frac + (9 - t); same closed form as k_scope_resample's fast weights
		const int j = 9 - t;
		const int i = ib - j;
		const double d = frac + double(j);
		if (i < 0 || i >= 21) return 0.0;
		if (d == 0.0) return 1.0;
		const double S1 = sinpi(frac);
		double sa, ca;
		sincospi(frac * 0.1, &sa, &ca);
		const int aj = j < 0 ? -j : j;
		const double cj = kCos10[aj], sj = j < 0 ? -kSin10[aj] : kSin10[aj];
		const double invPi2x10 = 10.0 / (3.1415926535897932384626433832795 * 3.1415926535897932384626433832795);
		return ((aj & 1) ? -S1 : S1) * (sa * cj + ca * sj) * invPi2x10 / (d * d);
	}

	__device__ __noinline__ double resampleVertexDirect(const ResampleArgs& a, const float* src, long long base, double delta)
	{
		const double x = 10.0 + delta;
		const double fx = floor(x);
		const int ib = int(fx);
		double acc = 0.0;
		for (int t = 0; t < 20; ++t)
		{
			const int i = ib - 9 + t;
			const long long idx = base + i;
			if (i >= 0 && i < 21 && idx >= 0 && idx < (long long)a.n)
			{
				float l, rr;
				loadLR(src, a.layout, a.n, uint64_t(idx), l, rr);
				acc += double(oscChannel(a.channel, l, rr)) * lanczosWeight10(x - fx, ib, t);
			}
		}
		return acc;
	}

	// per stream: the R x 20 steady-state weights (+ delta and kernel index of each phase), evaluated once per call
	// instead of once per CTA - the sinpi / sincospi / divide chain is ~20x the cost of using the weights
	struct PolyPhase { double w[20]; double delta; long long ib; }; // 176 bytes
	__global__ void k_scope_poly_weights(ResampleArgs a, int R, PolyPhase* __restrict__ table)
	{
		const uint32_t s = blockIdx.x;
		const double samplePos0 = a.samplePos[s];
		const double fl0 = floor(samplePos0);
		for (int e = threadIdx.x; e < R * 20; e += blockDim.x)
		{
			const int r = e / 20, t = e - r * 20;
			// steady-state delta of phase r: taken at an output far enough along that the window slides (q = 2)
			const double currentSample = fl0 + double(2 * R + r) * a.spp;
			const double d0 = currentSample - samplePos0;
			const double k = d0 > 1.0 ? ceil(d0 - 1.0) : 0.0;
			const double delta = currentSample - (samplePos0 + k);
			const double x = 10.0 + delta;
			const double fx = floor(x);
			PolyPhase& ph = table[uint64_t(s) * R + r];
			if (t == 0) { ph.delta = delta; ph.ib = (long long)fx; }
			ph.w[t] = lanczosWeight10(x - fx, int(fx), t);
		}
	}

	template <int LAYOUT, int CHANNEL>
	__global__ void __launch_bounds__(256, SGZ_POLY_MINB) k_scope_resample_poly(ResampleArgs a, int R, const PolyPhase* __restrict__ table)
	{
		extern __shared__ double xsamples[]; // kPolyBlockVertices / R + 32
		const uint32_t s = blockIdx.y;
		const double samplePos0 = a.samplePos[s];
		const double fl0 = floor(samplePos0);
		const float* src = a.audio + uint64_t(s) * a.stride;
		const uint64_t mBlock = uint64_t(blockIdx.x) * kPolyBlockVertices;
		const uint64_t mLast = min(mBlock + kPolyBlockVertices, a.count) - 1;
		// window slide k of a vertex (closed form of the reference's while (delta > 1) loop), nondecreasing in m
		auto slideOf = [&](uint64_t m, double& delta) -> long long
		{
			const double currentSample = fl0 + double(m) * a.spp;
			const double d0 = currentSample - samplePos0;
			const double k = d0 > 1.0 ? ceil(d0 - 1.0) : 0.0;
			delta = currentSample - (samplePos0 + k);
			return (long long)k;
		};
		// every global read of the block is issued up front so that their latencies overlap: this thread's phase
		// weights, the x coordinates of its vertices, then (below) the block's samples
		const int r = int(uint32_t(mBlock + threadIdx.x) & uint32_t(R - 1));
		const PolyPhase& ph = table[uint64_t(s) * R + r];
		double w[20];
#pragma unroll
		for (int t = 0; t < 20; ++t) w[t] = ph.w[t];
		const int ib = int(ph.ib);
		const double deltaR = ph.delta;
		// the samples this block's vertices can touch, converted to double ONCE (the float -> double conversion
		// unit is 8x narrower than the fp64 FMA pipe, and every sample feeds ~20 R taps); outside the buffer = 0
		double unused;
		const long long origin = (long long)a.n - (long long)fl0 - 10; // index of kernel[0] before any slide
		const long long lo = origin + slideOf(mBlock, unused) + 1;     // table taps start at kernel index ib - 9 >= 1
		const long long hi = origin + slideOf(mLast, unused) + 21;
		const int span = int(hi - lo + 1);                             // <= kPolyBlockVertices / R + 23
		for (int e = threadIdx.x; e < span; e += blockDim.x)
		{
			const long long idx = lo + e;
			float v = 0.f;
			if (idx >= 0 && idx < (long long)a.n)
			{
				float l, rr;
				loadLR(src, LAYOUT, a.n, uint64_t(idx), l, rr);
				v = oscChannel(CHANNEL, l, rr);
			}
			xsamples[e] = double(v);
		}
		__syncthreads();
		// vertices m = mBlock + tid + 256 j: one phase per thread (256 is a multiple of R), so its 20 weights live in
		// registers; neighbouring lanes read the same or neighbouring samples (broadcast / conflict-free)
		// three sweeps over the thread's vertices so that independent work interleaves (the kernel is latency-bound,
		// not throughput-bound): window slides, then the 20-tap sums with one accumulator chain per vertex, then stores
		for (int g = 0; g < kPolyGroups; ++g)
		{
			const uint64_t mGroup = mBlock + uint64_t(g) * (256 * kPolyVerticesPerThread);
			if (mGroup >= a.count) break;
			float xv[kPolyVerticesPerThread];
#pragma unroll
			for (int it = 0; it < kPolyVerticesPerThread; ++it)
			{
				const uint64_t m = mGroup + threadIdx.x + uint64_t(it) * 256;
				xv[it] = m < a.count ? a.xs[m] : 0.f;
			}
			int offs[kPolyVerticesPerThread];   // sample index of tap 0 relative to xsamples; < 0: not a table vertex
			double accs[kPolyVerticesPerThread];
#pragma unroll
			for (int it = 0; it < kPolyVerticesPerThread; ++it)
			{
				const uint64_t m = mGroup + threadIdx.x + uint64_t(it) * 256;
				double delta;
				const long long base = origin + slideOf(m, delta);
				// samplePos + k rounds differently as k grows, so the deltas of one phase agree to ~1e-12 rather
				// than exactly; the weights are smooth in delta (|dw/d delta| < 4), far inside the float32 output ulp
				const bool table = m < a.count && fabs(delta - deltaR) < 1e-9;
				offs[it] = table ? int(base + ib - 9 - lo) : -1;
				accs[it] = 0.0;
			}
			// ib is 10, or 11 when the vertex sits exactly on a sample (delta == 1): then the last tap falls outside the
			// 21-sample kernel and is skipped, as the reference's bounds check does
			const int taps = ib == 10 ? 20 : 19;
#pragma unroll
			for (int t = 0; t < 20; ++t)
			{
				const double wt = w[t];
#pragma unroll
				for (int it = 0; it < kPolyVerticesPerThread; ++it)
					if (offs[it] >= 0 && t < taps) accs[it] += xsamples[offs[it] + t] * wt;
			}
#pragma unroll
			for (int it = 0; it < kPolyVerticesPerThread; ++it)
			{
				const uint64_t m = mGroup + threadIdx.x + uint64_t(it) * 256;
				if (m >= a.count) break;
				double acc = accs[it];
				if (offs[it] < 0)
				{
					double delta;
					const long long base = origin + slideOf(m, delta);
					acc = resampleVertexDirect(a, src, base, delta); // the first outputs, before the window starts sliding
				}
				reinterpret_cast<float2*>(a.out)[uint64_t(s) * a.count + m] = make_float2(xv[it], float(acc));
			}
	
		}
	}

	// ------------------------------------------------------------------------------------------
	// Oscilloscope: stream triggers.  One warp per stream, 32 samples per step.
	// ------------------------------------------------------------------------------------------
	struct TriggerArgs
	{
		const float* audio; int layout; uint64_t n; uint64_t stride; int channel; uint32_t nStreams;
		double threshold, hysteresis; uint64_t steadyClock;
		sgz_scope_trigger_state* states; uint64_t* peaks; uint32_t* counts; uint32_t maxPeaks;
	};

	// ZeroCrossingProcessor::process (StreamPreprocessing.h:331-347).  Per sample: a rising zero crossing arms the
	// trigger and remembers where (holding = true, origin = position); a sample above the threshold while armed
	// emits that origin and disarms.  The state entering a stretch of samples only matters up to the stretch's first
	// event: a crossing overwrites it, and an above-threshold sample that comes before any crossing emits the
	// INCOMING origin (if armed) and disarms.  So a stream is cut into kTrigWarps contiguous chunks, one warp each:
	//   pass A  every warp runs the automaton over its chunk from the disarmed state (coalesced 32-sample steps,
	//           predicates by ballot, all 32 positions resolved by one carry-chain addition) and records: emits inside the chunk, whether an
	//           above-threshold sample precedes every crossing, whether there was any event, and its final state;
	//   pass B  one lane chains the kTrigWarps summaries (16 steps) into each chunk's true incoming state and the
	//           output offset of its first emit;
	//   pass C  every warp replays its chunk (now from L2) from the true incoming state and writes its origins.
	// Positions come out in ascending order and bit-identical to the sequential scan.
	constexpr int kTrigWarps = 16;
	constexpr int kTrigRows = 8; // 32-sample rows fetched per trip
	struct TrigSummary { uint32_t emits; int aboveFirst, anyEvent, holdingOut; uint64_t originOut; };

	template<bool WRITE>
	__device__ __forceinline__ void trigZeroChunk(const TriggerArgs& a, const float* base, uint64_t begin, uint64_t end, float prev, float aboveFrom, bool holding,
		uint64_t origin, uint64_t* out, uint32_t outPos, int lane, TrigSummary& sum, float& lastSample)
	{
		uint32_t emits = 0;
		int aboveFirst = 0, anyEvent = 0;
		// the rows of a trip are fetched together, one trip ahead of the scan, so that memory latency overlaps both
		// itself and the scan (which is a handful of integer operations per row; what it waits for is memory)
		auto fetch = [&](uint64_t n1, float (&dst)[kTrigRows])
		{
#pragma unroll
			for (int k = 0; k < kTrigRows; ++k)
			{
				const uint64_t i = n1 + 32 * k + lane;
				dst[k] = 0.f;
				if (i < end) { float l, r; loadLR(base, a.layout, a.n, i, l, r); dst[k] = oscChannel(a.channel, l, r); }
			}
		};
		float xs[kTrigRows], ahead[kTrigRows];
		fetch(begin, ahead);
		for (uint64_t n1 = begin; n1 < end; n1 += 32 * kTrigRows)
		{
#pragma unroll
		for (int k = 0; k < kTrigRows; ++k) xs[k] = ahead[k];
		if (n1 + 32 * kTrigRows < end) fetch(n1 + 32 * kTrigRows, ahead);
#pragma unroll
		for (int k = 0; k < kTrigRows; ++k)
		{
			const uint64_t n0 = n1 + 32 * k;
			if (n0 >= end) break;
			const bool valid = n0 + lane < end;
			// float comparisons that decide exactly like the reference's double ones: the signs are the same, and
			// sample > threshold <=> sample >= aboveFrom, the smallest float strictly greater than the threshold
			const float sample = xs[k];
			float before = __shfl_up_sync(0xffffffffu, sample, 1);
			if (lane == 0) before = prev;
			const uint32_t cross = __ballot_sync(0xffffffffu, valid && sample > 0.f && before < 0.f);
			const uint32_t above = __ballot_sync(0xffffffffu, valid && sample >= aboveFrom);
			uint32_t todo = cross | above;
			if (todo && !anyEvent)
			{
				// first event of the chunk: is it an above-threshold sample with no crossing at or before it?
				const int j = __ffs(todo) - 1;
				aboveFirst = ((above >> j) & 1u) && !((cross >> j) & 1u);
				anyEvent = 1;
			}
			if (todo)
			{
				// the armed flag through the 32 samples is a carry chain: a crossing that is not also above the threshold
				// generates it, an above-threshold sample kills it, every other sample propagates it - so one 64-bit
				// addition resolves all 32 positions at once (bit j of `carries` = armed after sample j - 1)
				const uint32_t gen = cross & ~above, prop = ~(cross | above);
				const uint64_t X = uint64_t(gen | prop), Y = uint64_t(gen);
				const uint64_t carries = (X + Y + (holding ? 1u : 0u)) ^ X ^ Y;
				const uint32_t fires = above & (cross | uint32_t(carries)); // armed (or arming right now) and above
				const int nFires = __popc(fires);
				if (WRITE && ((fires >> lane) & 1u))
				{
					// the lane of a firing sample reports it: its rank among the firing samples gives the slot, the
					// latest crossing at or before it the origin
					const uint32_t slot = outPos + emits + __popc(fires & ((1u << lane) - 1u));
					const uint32_t earlier = cross & (0xffffffffu >> (31 - lane));
					if (slot < a.maxPeaks) out[slot] = earlier ? a.steadyClock + n0 + (31 - __clz(earlier)) : origin;
				}
				emits += nFires;
				if (cross) origin = a.steadyClock + n0 + (31 - __clz(cross));
				holding = ((carries >> 32) & 1u) != 0;
			}
			const int last = int(min(uint64_t(32), end - n0)) - 1;
			prev = __shfl_sync(0xffffffffu, sample, last);
		}
		}
		sum.emits = emits; sum.aboveFirst = aboveFirst; sum.anyEvent = anyEvent; sum.holdingOut = holding ? 1 : 0; sum.originOut = origin;
		lastSample = prev;
	}

	__global__ void __launch_bounds__(32 * kTrigWarps) k_scope_trigger_zero(TriggerArgs a)
	{
		__shared__ TrigSummary sums[kTrigWarps];
		__shared__ int holdIn[kTrigWarps];
		__shared__ uint64_t originIn[kTrigWarps];
		__shared__ uint32_t offset[kTrigWarps + 1];
		const uint32_t stream = blockIdx.x;
		const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
		const float* base = a.audio + uint64_t(stream) * a.stride;
		const sgz_scope_trigger_state st = a.states[stream];
		// chunk bounds: multiples of 32 samples so every step is a whole aligned row
		const uint64_t per = ((a.n + kTrigWarps - 1) / kTrigWarps + 31) & ~uint64_t(31);
		const uint64_t begin = min(uint64_t(warp) * per, a.n), end = min(begin + per, a.n);
		// the sample before the chunk (the crossing test of its first sample): the carried state for chunk 0 (a float
		// widened to double by the previous call; any other double keeps at least its sign through the conversion)
		float prev = float(st.state);
		if (st.state != 0.0 && prev == 0.f) prev = st.state < 0 ? -1.f : 1.f;
		if (begin > 0 && begin < a.n) { float l, r; loadLR(base, a.layout, a.n, begin - 1, l, r); prev = oscChannel(a.channel, l, r); }
		// smallest float strictly greater than the (double) threshold
		float aboveFrom = float(a.threshold);
		if (!(double(aboveFrom) > a.threshold)) aboveFrom = nextafterf(aboveFrom, INFINITY);
		TrigSummary mine;
		float lastSample = prev;
		trigZeroChunk<false>(a, base, begin, end, prev, aboveFrom, false, 0, nullptr, 0, lane, mine, lastSample);
		if (lane == 0) sums[warp] = mine;
		__syncthreads();
		if (threadIdx.x == 0)
		{
			int holding = st.is_peak_holding != 0;
			uint64_t origin = st.cross_origin;
			uint32_t total = 0;
			for (int w = 0; w < kTrigWarps; ++w)
			{
				holdIn[w] = holding; originIn[w] = origin; offset[w] = total;
				const TrigSummary& c = sums[w];
				total += c.emits + ((c.aboveFirst && holding) ? 1u : 0u);
				if (c.anyEvent) { holding = c.holdingOut; origin = c.holdingOut || c.emits ? c.originOut : origin; }
			}
			offset[kTrigWarps] = total;
			a.counts[stream] = total;
			// .state (the last sample of the block) is written by the warp that owns the end of the stream, below
			a.states[stream].is_peak_holding = holding;
			a.states[stream].cross_origin = origin;
		}
		__syncthreads();
		// replay with the true incoming state only where something is written
		const bool extra = sums[warp].aboveFirst && holdIn[warp];
		if ((sums[warp].emits || extra) && offset[warp] < a.maxPeaks)
		{
			TrigSummary again; float ls;
			trigZeroChunk<true>(a, base, begin, end, prev, aboveFrom, holdIn[warp] != 0, originIn[warp], a.peaks + uint64_t(stream) * a.maxPeaks, offset[warp], lane, again, ls);
		}
		if (lane == 0 && end == a.n && begin < end) a.states[stream].state = double(lastSample);
	}

	// PeakHoldProcessor::process (StreamPreprocessing.h:284-311): a data-dependent double recurrence;
	// lanes fetch, lane order is replayed sequentially
	__global__ void k_scope_trigger_peak(TriggerArgs a)
	{
		const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
		if (warp >= a.nStreams) return;
		const float* base = a.audio + uint64_t(warp) * a.stride;
		sgz_scope_trigger_state st = a.states[warp];
		double state = st.state;
		bool holding = st.is_peak_holding != 0;
		uint32_t count = 0;
		const double thr2 = __dmul_rn(a.threshold, a.threshold);
		for (uint64_t n0 = 0; n0 < a.n; n0 += 32)
		{
			float x = 0.f;
			if (n0 + lane < a.n) { float l, r; loadLR(base, a.layout, a.n, n0 + lane, l, r); x = oscChannel(a.channel, l, r); }
			const double sq = __dmul_rn(double(x), double(x));
			const int cnt = int(min(uint64_t(32), a.n - n0));
			for (int j = 0; j < cnt; ++j)
			{
				const double sample = __shfl_sync(0xffffffffu, sq, j);
				const double delta = __dsub_rn(sample, state);
				if (delta < 0)
				{
					state = __dmul_rn(state, 0.9999);
					state = fmax(thr2, state);
					if (holding)
					{
						if (lane == 0 && count < a.maxPeaks) a.peaks[uint64_t(warp) * a.maxPeaks + count] = a.steadyClock + n0 + j - 1;
						++count;
						holding = false;
					}
				}
				else
				{
					if (delta > __dmul_rn(a.hysteresis, state)) holding = true;
					state = sample;
				}
			}
		}
		if (lane == 0)
		{
			st.state = state; st.is_peak_holding = holding ? 1 : 0;
			a.states[warp] = st;
			a.counts[warp] = count;
		}
	}

	// ------------------------------------------------------------------------------------------
	// Oscilloscope: spectral period search.  One CTA per stream: 8192-point float64 FFT in shared
	// memory (128 KB), the sequential pick by one thread, then a second kernel for the Goertzel sum.
	// ------------------------------------------------------------------------------------------
	constexpr int kLookahead = 8192;

	__device__ __forceinline__ double2 dmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
	__device__ __forceinline__ double dabs(double2 a) { return hypot(a.x, a.y); }

	struct PeriodPickArgs
	{
		const float* audio; int layout; uint64_t n; uint64_t stride; int channel;
		long long firstSample; // n - max(ceil(effectiveWindowSize), 8192)
		double threshold, hysteresis;
		sgz_scope_bin_record* picked; // [S]
	};

	__device__ double quadDelta(const double2* tb, int w)
	{
		const double2 x0 = tb[w], x1 = tb[w + 1], xm1 = tb[w == 0 ? 1 : w - 1];
		const double2 denom = make_double2(x0.x * 2.0 - xm1.x - x1.x, x0.y * 2.0 - xm1.y - x1.y);
		if ((denom.x + denom.y) == 0) return 0;
		const double2 num = make_double2(xm1.x - x1.x, xm1.y - x1.y);
		// Re(num / denom)
		return (num.x * denom.x + num.y * denom.y) / (denom.x * denom.x + denom.y * denom.y);
	}

	// W_8192^k = exp(-2 pi i k / 8192), k < 4096, for the fp64 FFT below: filled once per device by k_scope_period_twiddles
	// (sincospi of the exact argument, so every stage's W_(2 half)^j = table[j (4096 / half)] has the bits a direct
	// evaluation would give)
	static __device__ double2 g_periodTwiddle[kLookahead / 2];

	__global__ void k_scope_period_twiddles()
	{
		const int k = blockIdx.x * blockDim.x + threadIdx.x;
		if (k >= kLookahead / 2) return;
		double sn, cs;
		sincospi(-double(k) / double(kLookahead / 2), &sn, &cs);
		g_periodTwiddle[k] = make_double2(cs, sn);
	}

	// calculateFundamentalPeriod (OscilloscopeDSP.inl:61-223) up to the pick: fp64 8192-point FFT of one stream in shared
	// memory (radix 2, twiddles from the table), |X| of the 4095 candidate bins in parallel, then the reference's scan
	// (:128-184).  The scan carries `max` from bin to bin, but a bin only matters if it passes the entry test against the
	// CURRENT max (`invHysteresis |X_i| > 2 max`); warp 0 tests 32 bins per step by ballot and lane 0 runs the reference's
	// decision on the ones that pass, in order, re-testing each against the max as it stands.
	__global__ void __launch_bounds__(512) k_scope_period_pick(PeriodPickArgs a)
	{
		extern __shared__ double2 tb[];
		const uint32_t s = blockIdx.x;
		const float* base = a.audio + uint64_t(s) * a.stride;
		constexpr int N = kLookahead, bits = 13;
		for (int i = threadIdx.x; i < N; i += blockDim.x)
		{
			const long long idx = a.firstSample + i;
			float v = 0.f;
			if (idx >= 0 && idx < (long long)a.n) { float l, r; loadLR(base, a.layout, a.n, uint64_t(idx), l, r); v = oscChannel(a.channel, l, r); }
			tb[__brev(unsigned(i)) >> (32 - bits)] = make_double2(double(v), 0.0);
		}
		__syncthreads();
		for (int half = 1; half < N; half <<= 1)
		{
			const int step = (N >> 1) / half;
			for (int b = threadIdx.x; b < (N >> 1); b += blockDim.x)
			{
				const int j = b & (half - 1);
				const int i0 = ((b - j) << 1) + j;
				const double2 w = g_periodTwiddle[j * step]; // W_(2 half)^j
				const double2 u = tb[i0];
				const double2 v = dmul(tb[i0 + half], w);
				tb[i0] = make_double2(u.x + v.x, u.y + v.y);
				tb[i0 + half] = make_double2(u.x - v.x, u.y - v.y);
			}
			__syncthreads();
		}
		// |X_i| of the bins the scan looks at, into the upper half of the buffer (bins >= N / 2 are never read)
		double* mag = reinterpret_cast<double*>(tb + (N >> 1));
		for (int i = threadIdx.x; i < (N >> 1); i += blockDim.x) mag[i] = dabs(tb[i]);
		__syncthreads();
		if (threadIdx.x >= 32) return;
		const int lane = threadIdx.x;
		const double quarterSemitone = pow(2.0, 0.25 / 12.0) - 1;
		const double invHysteresis = 1 - a.hysteresis;
		uint64_t mIndex = 1; double mValue = fmax(a.threshold * N / 6.0, mag[1]), mOffset = quadDelta(tb, 1);
		for (int i0 = 2; i0 < (N >> 1); i0 += 32)
		{
			const int i = i0 + lane;
			const double cMine = i < (N >> 1) ? mag[i] : 0.0;
			// mValue only grows, so a bin that fails against the max at the start of the step fails against any later one
			unsigned pending = __ballot_sync(0xffffffffu, invHysteresis * cMine > mValue * 2);
			while (pending)
			{
				const int l = __ffs(pending) - 1;
				pending &= pending - 1;
				if (lane == 0)
				{
					const int c = i0 + l;
					const double cValue = mag[c];
					if (invHysteresis * cValue > mValue * 2)
					{
						const double mOmega = double(mIndex) + mOffset;
						if (mOmega > 0)
						{
							const double cOffset = quadDelta(tb, c);
							const double factor = (double(c) + cOffset) / mOmega;
							const double sensivity = cValue / mValue;
							bool take = false;
							if (invHysteresis * sensivity > 20) take = true;
							else if (fabs(1 - factor) < invHysteresis * quarterSemitone) take = true;
							else
							{
								const double multipleDeviation = fabs(factor - floor(factor + 0.5));
								if (invHysteresis * fabs(multipleDeviation) > quarterSemitone) take = true;
							}
							if (take) { mIndex = c; mValue = cValue; mOffset = cOffset; }
						}
						else
						{
							mIndex = c; mValue = cValue; mOffset = quadDelta(tb, c);
						}
					}
				}
			}
			mValue = __shfl_sync(0xffffffffu, mValue, 0); // the next step's entry test uses lane 0's max
		}
		if (lane == 0) { a.picked[s].index = mIndex; a.picked[s].value = mValue; a.picked[s].offset = mOffset; }
	}

	struct GoertzelArgs
	{
		const float* audio; int layout; uint64_t n; uint64_t stride; int channel;
		const long long* firstSample; // [S]
		const double* radians;        // [S]
		double2* z;                   // [S]
	};

	// cpl::dsp::goertzel(buf, N, w) = sum x[n] exp(-i w n) over the 8192 newest samples (:271-277)
	__global__ void __launch_bounds__(256) k_scope_goertzel(GoertzelArgs a)
	{
		__shared__ double2 red[256];
		const uint32_t s = blockIdx.x;
		const float* base = a.audio + uint64_t(s) * a.stride;
		const double w = a.radians[s];
		const long long first = a.firstSample[s];
		double2 acc = make_double2(0.0, 0.0);
		for (int i = threadIdx.x; i < kLookahead; i += blockDim.x)
		{
			const long long idx = first + i;
			float v = 0.f;
			if (idx >= 0 && idx < (long long)a.n) { float l, r; loadLR(base, a.layout, a.n, uint64_t(idx), l, r); v = oscChannel(a.channel, l, r); }
			double sn, cs;
			sincos(w * double(i), &sn, &cs);
			acc.x += double(v) * cs; acc.y -= double(v) * sn;
		}
		red[threadIdx.x] = acc;
		__syncthreads();
		for (int stride = 128; stride > 0; stride >>= 1)
		{
			if (threadIdx.x < stride) { red[threadIdx.x].x += red[threadIdx.x + stride].x; red[threadIdx.x].y += red[threadIdx.x + stride].y; }
			__syncthreads();
		}
		if (threadIdx.x == 0) a.z[s] = red[0];
	}
#include "sgz_colour.cuh"
}

using namespace sgz;

#define SGZS_GUARD(ctx)                                                                  \
	if (!(ctx)) return fail(SGZ_EINVAL, "null context");                                 \
	CtxGuard g__((ctx)->device);                                                         \
	if (!g__.ok) return fail(SGZ_ECUDA, "cudaSetDevice failed")

extern "C"
{
	int sgz_context_create(int device, sgz_context** out)
	{
		if (!out) return fail(SGZ_EINVAL, "null argument");
		*out = nullptr;
		int count = 0;
		if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return fail(SGZ_ECUDA, "no CUDA device (this library has no CPU fallback)");
		if (device < 0 || device >= count) return fail(SGZ_EINVAL, "device %d out of range (%d devices)", device, count);
		sgz_context* c = new (std::nothrow) sgz_context();
		if (!c) return fail(SGZ_ENOMEM, "out of host memory");
		c->device = device;
		CtxGuard g(device);
		if (!g.ok || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return fail(SGZ_ECUDA, "cannot create a stream on device %d", device); }
		c->ownStream = true;
		*out = c;
		return SGZ_OK;
	}

	void sgz_context_destroy(sgz_context* c)
	{
		if (!c) return;
		{
			CtxGuard g(c->device);
			if (c->stream) cudaStreamSynchronize(c->stream);
			for (void* p : c->stage) if (p) cudaFree(p);
			if (c->xTable) cudaFree(c->xTable);
			if (c->ownStream && c->stream) cudaStreamDestroy(c->stream);
		}
		delete c;
	}

	int sgz_context_set_cuda_stream(sgz_context* c, void* cuda_stream)
	{
		SGZS_GUARD(c);
		if (c->stream) cudaStreamSynchronize(c->stream);
		if (c->ownStream && c->stream) cudaStreamDestroy(c->stream);
		c->stream = static_cast<cudaStream_t>(cuda_stream);
		c->ownStream = false;
		return SGZ_OK;
	}

	int sgz_context_synchronize(sgz_context* c)
	{
		SGZS_GUARD(c);
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		return SGZ_OK;
	}

	static int vscopeCheck(const sgz_vscope_config* cfg)
	{
		if (!cfg || cfg->abi_version != SGZ_ABI_VERSION) return fail(SGZ_EINVAL, "bad vectorscope config");
		if (cfg->width == 0 || cfg->height == 0 || cfg->width > 16384 || cfg->height > 16384) return fail(SGZ_EINVAL, "bad histogram size");
		return SGZ_OK;
	}

	int sgz_vscope_vertices(sgz_context* c, const sgz_vscope_config* cfg, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, float* xyz, int out_mem)
	{
		int rc = vscopeCheck(cfg); if (rc != SGZ_OK) return rc;
		if (!xyz) return fail(SGZ_EINVAL, "null output");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		if (n_samples == 0) return SGZ_OK;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		const size_t bytes = size_t(n_streams) * n_samples * 3 * sizeof(float);
		float* dOut = xyz;
		if (out_mem == SGZ_MEM_HOST) { void* p = nullptr; if ((rc = ensure(c, 1, bytes, &p)) != SGZ_OK) return rc; dOut = static_cast<float*>(p); }
		const dim3 block(256), grid(unsigned((n_samples + 255) / 256), n_streams);
		k_vscope_vertices<<<grid, block, 0, c->stream>>>(*cfg, d, layout, n_samples, stream_stride, dOut);
		SGZS_CUDA(cudaGetLastError());
		if (out_mem == SGZ_MEM_HOST)
		{
			SGZS_CUDA(cudaMemcpyAsync(xyz, dOut, bytes, cudaMemcpyDeviceToHost, c->stream));
			SGZS_CUDA(cudaStreamSynchronize(c->stream));
		}
		return SGZ_OK;
	}

	int sgz_vscope_histogram(sgz_context* c, const sgz_vscope_config* cfg, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, uint32_t* hist, int out_mem)
	{
		int rc = vscopeCheck(cfg); if (rc != SGZ_OK) return rc;
		if (!hist) return fail(SGZ_EINVAL, "null output");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if (n_samples && (rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		const size_t bytes = size_t(n_streams) * cfg->width * cfg->height * sizeof(uint32_t);
		uint32_t* dOut = hist;
		if (out_mem == SGZ_MEM_HOST) { void* p = nullptr; if ((rc = ensure(c, 1, bytes, &p)) != SGZ_OK) return rc; dOut = static_cast<uint32_t*>(p); }
		SGZS_CUDA(cudaMemsetAsync(dOut, 0, bytes, c->stream));
		if (n_samples)
		{
			// enough CTAs per stream to fill the machine when there are few streams, grid-stride inside
			unsigned perStream = unsigned(std::min<uint64_t>((n_samples + 255) / 256, std::max<uint64_t>(1, (148ull * 8 + n_streams - 1) / n_streams)));
			k_vscope_histogram<<<dim3(perStream, n_streams), 256, 0, c->stream>>>(*cfg, d, layout, n_samples, stream_stride, dOut);
			SGZS_CUDA(cudaGetLastError());
		}
		if (out_mem == SGZ_MEM_HOST)
		{
			SGZS_CUDA(cudaMemcpyAsync(hist, dOut, bytes, cudaMemcpyDeviceToHost, c->stream));
			SGZS_CUDA(cudaStreamSynchronize(c->stream));
		}
		return SGZ_OK;
	}

	int sgz_vscope_meters(sgz_context* c, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
		float envelope_coeff, float stereo_coeff, float second_stereo_filter_speed, sgz_vscope_filters* filters)
	{
		int rc;
		if (!filters) return fail(SGZ_EINVAL, "null filters");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		if (n_samples == 0) return SGZ_OK;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		// host slot order: envelope[2], balance[slow][L,R], balance[fast][L,R], phase[slow], phase[fast] = struct order
		void* p = nullptr; if ((rc = ensure(c, 2, size_t(n_streams) * 8 * sizeof(float), &p)) != SGZ_OK) return rc;
		SGZS_CUDA(cudaMemcpyAsync(p, filters, size_t(n_streams) * 8 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
		const float pole1 = static_cast<float>(std::pow(stereo_coeff, second_stereo_filter_speed)); // Vectorscope.cpp:279
		const unsigned warpsPerBlock = 4;
		k_vscope_meters<<<(n_streams + warpsPerBlock - 1) / warpsPerBlock, warpsPerBlock * 32, 0, c->stream>>>(d, layout, n_samples, stream_stride, n_streams,
			envelope_coeff, stereo_coeff, pole1, static_cast<float*>(p));
		SGZS_CUDA(cudaGetLastError());
		SGZS_CUDA(cudaMemcpyAsync(filters, p, size_t(n_streams) * 8 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		return SGZ_OK;
	}

	uint64_t sgz_scope_resample_count(const sgz_scope_resample_config* cfg)
	{
		if (!cfg || !(cfg->inc > 0) || !(cfg->right >= cfg->left)) return 0;
		// the do-while at OscilloscopeRendering.cpp:863-886, replayed so the count is exact
		uint64_t count = 0;
		double unitSpacePos = cfg->left;
		do { ++count; unitSpacePos += cfg->inc; } while (unitSpacePos < (cfg->right + cfg->inc) && count < (1ull << 32));
		return count;
	}

	int sgz_scope_resample(sgz_context* c, const sgz_scope_resample_config* cfg, const double* sample_pos, const float* audio, int mem, int layout,
		uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride, float* out_xy, int out_mem)
	{
		int rc;
		if (!cfg || cfg->abi_version != SGZ_ABI_VERSION || !sample_pos || !out_xy) return fail(SGZ_EINVAL, "bad resample arguments");
		if (cfg->channel < SGZ_OSC_LEFT || cfg->channel > SGZ_OSC_SIDE) return fail(SGZ_EINVAL, "bad channel");
		if (!(cfg->samples_per_pixel > 0)) return fail(SGZ_EINVAL, "samples_per_pixel must be positive");
		if (!c) return fail(SGZ_EINVAL, "null context");
		const bool sameView = c->xTable && c->xCount && c->xKey[0] == cfg->left && c->xKey[1] == cfg->right && c->xKey[2] == cfg->inc;
		const uint64_t count = sameView ? c->xCount : sgz_scope_resample_count(cfg);
		if (count == 0) return fail(SGZ_EINVAL, "empty output range");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		void* dPos = nullptr; if ((rc = ensure(c, 2, size_t(n_streams) * sizeof(double), &dPos)) != SGZ_OK) return rc;
		SGZS_CUDA(cudaMemcpyAsync(dPos, sample_pos, size_t(n_streams) * sizeof(double), cudaMemcpyHostToDevice, c->stream));
		const size_t bytes = size_t(n_streams) * count * 2 * sizeof(float);
		float* dOut = out_xy;
		if (out_mem == SGZ_MEM_HOST) { void* p = nullptr; if ((rc = ensure(c, 1, bytes, &p)) != SGZ_OK) return rc; dOut = static_cast<float*>(p); }
		if (reinterpret_cast<uintptr_t>(dOut) & 7) return fail(SGZ_EINVAL, "out_xy must be 8-byte aligned");
		ResampleArgs a{};
		a.audio = d; a.layout = layout; a.n = n_samples; a.stride = stream_stride; a.channel = cfg->channel;
		if (!(c->xTable && c->xCount == count && c->xKey[0] == cfg->left && c->xKey[1] == cfg->right && c->xKey[2] == cfg->inc))
		{
			std::vector<float> xs(count);
			double unitSpacePos = cfg->left;
			for (uint64_t m = 0; m < count; ++m) { xs[m] = static_cast<float>(unitSpacePos); unitSpacePos += cfg->inc; }
			if (c->xCapacity < count)
			{
				if (c->xTable) { cudaFree(c->xTable); c->xTable = nullptr; c->xCapacity = 0; }
				SGZS_CUDA(cudaMalloc(&c->xTable, size_t(count) * sizeof(float)));
				c->xCapacity = count;
			}
			c->xCount = 0;
			SGZS_CUDA(cudaMemcpyAsync(c->xTable, xs.data(), size_t(count) * sizeof(float), cudaMemcpyHostToDevice, c->stream));
			SGZS_CUDA(cudaStreamSynchronize(c->stream)); // xs is a local
			c->xCount = count; c->xKey[0] = cfg->left; c->xKey[1] = cfg->right; c->xKey[2] = cfg->inc;
		}
		a.samplePos = static_cast<const double*>(dPos); a.spp = cfg->samples_per_pixel; a.count = count; a.xs = c->xTable; a.out = dOut;
		// integer oversampling ratio (a power of two keeps m * spp exact) -> polyphase weights
		const double ratio = 1.0 / cfg->samples_per_pixel;
		const int R = int(ratio);
		if (double(R) == ratio && R >= 1 && R <= kMaxPhases && (R & (R - 1)) == 0)
		{
			const dim3 grid(unsigned((count + kPolyBlockVertices - 1) / kPolyBlockVertices), n_streams);
			const size_t polySmem = (size_t(kPolyBlockVertices) / R + 32) * sizeof(double);
			void* tp = nullptr; if ((rc = ensure(c, 3, size_t(n_streams) * R * sizeof(PolyPhase), &tp)) != SGZ_OK) return rc;
			PolyPhase* table = static_cast<PolyPhase*>(tp);
			k_scope_poly_weights<<<n_streams, 320, 0, c->stream>>>(a, R, table);
#define SGZ_POLY(L, C) { if (polySmem > 48 * 1024) SGZS_CUDA(cudaFuncSetAttribute(k_scope_resample_poly<L, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(polySmem))); \
			k_scope_resample_poly<L, C><<<grid, 256, polySmem, c->stream>>>(a, R, table); }
#define SGZ_POLY_CH(L) \
			switch (cfg->channel) { case SGZ_OSC_LEFT: SGZ_POLY(L, SGZ_OSC_LEFT); break; case SGZ_OSC_RIGHT: SGZ_POLY(L, SGZ_OSC_RIGHT); break; \
				case SGZ_OSC_MID: SGZ_POLY(L, SGZ_OSC_MID); break; default: SGZ_POLY(L, SGZ_OSC_SIDE); break; }
			if (layout == SGZ_LAYOUT_INTERLEAVED) { SGZ_POLY_CH(SGZ_LAYOUT_INTERLEAVED) } else { SGZ_POLY_CH(SGZ_LAYOUT_PLANAR) }
#undef SGZ_POLY_CH
#undef SGZ_POLY
		}
		else
			k_scope_resample<<<dim3(unsigned((count + 127) / 128), n_streams), 128, 0, c->stream>>>(a);
		SGZS_CUDA(cudaGetLastError());
		if (out_mem == SGZ_MEM_HOST)
		{
			SGZS_CUDA(cudaMemcpyAsync(out_xy, dOut, bytes, cudaMemcpyDeviceToHost, c->stream));
			SGZS_CUDA(cudaStreamSynchronize(c->stream));
		}
		return SGZ_OK;
	}

	// the two trigger kernels on device-resident arrays (shared by the host-array and the device-array entry points)
	static int launchTrigger(sgz_context* c, int mode, int channel, const float* d, int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
		double threshold, double hysteresis, uint64_t steady_clock, sgz_scope_trigger_state* dStates, uint64_t* dPeaks, uint32_t* dCounts, uint32_t max_peaks)
	{
		TriggerArgs a{};
		a.audio = d; a.layout = layout; a.n = n_samples; a.stride = stream_stride; a.channel = channel; a.nStreams = n_streams;
		a.threshold = threshold; a.hysteresis = hysteresis; a.steadyClock = steady_clock; a.states = dStates; a.peaks = dPeaks; a.counts = dCounts; a.maxPeaks = max_peaks;
		const unsigned warpsPerBlock = 4, blocks = (n_streams + warpsPerBlock - 1) / warpsPerBlock;
		if (mode == 0) k_scope_trigger_zero<<<n_streams, 32 * kTrigWarps, 0, c->stream>>>(a);
		else k_scope_trigger_peak<<<blocks, warpsPerBlock * 32, 0, c->stream>>>(a);
		SGZS_CUDA(cudaGetLastError());
		return SGZ_OK;
	}

	// sgz_scope_trigger with every array on the device: nothing is copied and nothing waits - the call only enqueues the
	// kernel on the context's stream, so a device-resident pipeline keeps the kernel's rate
	int sgz_scope_trigger_device(sgz_context* c, int mode, int channel, const float* d_audio, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, double threshold, double hysteresis, uint64_t steady_clock, sgz_scope_trigger_state* d_states, uint64_t* d_peaks,
		uint32_t* d_counts, uint32_t max_peaks)
	{
		int rc;
		if (!d_states || !d_counts || (max_peaks && !d_peaks)) return fail(SGZ_EINVAL, "null argument");
		if (mode != 0 && mode != 1) return fail(SGZ_EINVAL, "mode must be 0 (zero crossing) or 1 (envelope hold)");
		if (channel < SGZ_OSC_LEFT || channel > SGZ_OSC_SIDE) return fail(SGZ_EINVAL, "bad channel");
		if ((rc = validateAudio(d_audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		return launchTrigger(c, mode, channel, d_audio, layout, n_streams, n_samples, stream_stride, threshold, hysteresis, steady_clock, d_states, d_peaks, d_counts, max_peaks);
	}

	int sgz_scope_trigger(sgz_context* c, int mode, int channel, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, double threshold, double hysteresis, uint64_t steady_clock, sgz_scope_trigger_state* states, uint64_t* peaks,
		uint32_t* counts, uint32_t max_peaks)
	{
		int rc;
		if (!states || !counts || (max_peaks && !peaks)) return fail(SGZ_EINVAL, "null argument");
		if (mode != 0 && mode != 1) return fail(SGZ_EINVAL, "mode must be 0 (zero crossing) or 1 (envelope hold)");
		if (channel < SGZ_OSC_LEFT || channel > SGZ_OSC_SIDE) return fail(SGZ_EINVAL, "bad channel");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if (n_samples && (rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		const size_t stBytes = size_t(n_streams) * sizeof(sgz_scope_trigger_state), pkBytes = size_t(n_streams) * std::max<uint32_t>(max_peaks, 1) * sizeof(uint64_t),
			cnBytes = size_t(n_streams) * sizeof(uint32_t);
		void* p = nullptr; if ((rc = ensure(c, 2, stBytes + pkBytes + cnBytes + 64, &p)) != SGZ_OK) return rc;
		auto* dStates = static_cast<sgz_scope_trigger_state*>(p);
		auto* dPeaks = reinterpret_cast<uint64_t*>(static_cast<char*>(p) + ((stBytes + 15) & ~size_t(15)));
		auto* dCounts = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(dPeaks) + pkBytes);
		SGZS_CUDA(cudaMemcpyAsync(dStates, states, stBytes, cudaMemcpyHostToDevice, c->stream));
		if ((rc = launchTrigger(c, mode, channel, d, layout, n_streams, n_samples, stream_stride, threshold, hysteresis, steady_clock, dStates, dPeaks, dCounts, max_peaks)) != SGZ_OK) return rc;
		SGZS_CUDA(cudaMemcpyAsync(states, dStates, stBytes, cudaMemcpyDeviceToHost, c->stream));
		if (max_peaks) SGZS_CUDA(cudaMemcpyAsync(peaks, dPeaks, size_t(n_streams) * max_peaks * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaMemcpyAsync(counts, dCounts, cnBytes, cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		return SGZ_OK;
	}

	int sgz_scope_period(sgz_context* c, const sgz_scope_period_params* prm, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, sgz_scope_period_state* states, sgz_scope_period_result* results)
	{
		int rc;
		if (!prm || prm->abi_version != SGZ_ABI_VERSION || !states || !results) return fail(SGZ_EINVAL, "bad period arguments");
		if (prm->channel < SGZ_OSC_LEFT || prm->channel > SGZ_OSC_SIDE) return fail(SGZ_EINVAL, "bad channel");
		if (!(prm->sample_rate > 0)) return fail(SGZ_EINVAL, "bad sample rate");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		constexpr size_t LookaheadSize = kLookahead, TransformSize = kLookahead;
		// --- calculateFundamentalPeriod: FFT + pick on the device ---
		const size_t recBytes = size_t(n_streams) * sizeof(sgz_scope_bin_record);
		const size_t auxBytes = size_t(n_streams) * (sizeof(long long) + sizeof(double) + sizeof(double2));
		void* p = nullptr; if ((rc = ensure(c, 2, recBytes + auxBytes + 64, &p)) != SGZ_OK) return rc;
		auto* dPicked = static_cast<sgz_scope_bin_record*>(p);
		auto* dFirst = reinterpret_cast<long long*>(static_cast<char*>(p) + ((recBytes + 15) & ~size_t(15)));
		auto* dRadians = reinterpret_cast<double*>(dFirst + n_streams);
		auto* dZ = reinterpret_cast<double2*>(dRadians + n_streams); // 16 n bytes past dFirst: 16-byte aligned
		PeriodPickArgs pa{};
		pa.audio = d; pa.layout = layout; pa.n = n_samples; pa.stride = stream_stride; pa.channel = prm->channel;
		const size_t offset1 = std::max<size_t>(static_cast<size_t>(std::ceil(prm->effective_window_size)), LookaheadSize); // :88
		pa.firstSample = (long long)n_samples - (long long)offset1;
		pa.threshold = prm->threshold; pa.hysteresis = prm->hysteresis; pa.picked = dPicked;
		const size_t smem = TransformSize * sizeof(double2);
		{
			// once per device: the opt-in shared-memory size and the twiddle table (stream-ordered before the first use)
			static std::atomic<uint64_t> ready{ 0 };
			const uint64_t bit = uint64_t(1) << (c->device & 63);
			if (!(ready.load(std::memory_order_acquire) & bit))
			{
				SGZS_CUDA(cudaFuncSetAttribute(k_scope_period_pick, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
				k_scope_period_twiddles<<<(kLookahead / 2 + 255) / 256, 256, 0, c->stream>>>();
				SGZS_CUDA(cudaGetLastError());
				SGZS_CUDA(cudaStreamSynchronize(c->stream));
				ready.fetch_or(bit, std::memory_order_release);
			}
		}
		k_scope_period_pick<<<n_streams, 512, smem, c->stream>>>(pa);
		SGZS_CUDA(cudaGetLastError());
		std::vector<sgz_scope_bin_record> picked(n_streams);
		SGZS_CUDA(cudaMemcpyAsync(picked.data(), dPicked, recBytes, cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		// --- median over the previous picks (:186-212) and the Goertzel parameters (:261-269), host side ---
		std::vector<long long> first(n_streams);
		std::vector<double> radians(n_streams), sampleDifference(n_streams);
		const double tau = 6.283185307179586476925286766559;
		for (uint32_t s = 0; s < n_streams; ++s)
		{
			sgz_scope_bin_record max = picked[s];
			sgz_scope_period_state& st = states[s];
			sgz_scope_bin_record local[8];
			std::memcpy(local, st.median, sizeof(local));
			st.median[st.median_pos & 7] = max;
			st.median_pos = (st.median_pos + 1) & 7;
			// std::nth_element leaves ties to the implementation; a stable order is the documented choice
			std::stable_sort(local, local + 8, [](const sgz_scope_bin_record& a, const sgz_scope_bin_record& b) { return a.index < b.index; });
			const sgz_scope_bin_record& old = local[4];
			const double omegaMax = double(max.index) + max.offset, omegaOld = double(old.index) + old.offset;
			if (old.index != uint64_t(-1) && std::abs(omegaMax - omegaOld) > 0.5) max = old;
			sgz_scope_period_result& r = results[s];
			r.record = max;
			const double omega = double(max.index) + max.offset;
			double fundamental = prm->sample_rate * omega / TransformSize;
			r.fundamental = fundamental = std::max(5.0, fundamental);
			r.cycle_samples = prm->sample_rate / fundamental;
			radians[s] = tau * omega / TransformSize;
			const double offsetReal = std::max<double>(double(LookaheadSize), prm->effective_window_size + r.cycle_samples);
			const size_t offset2 = static_cast<size_t>(std::ceil(offsetReal));
			sampleDifference[s] = double(offset2) - (prm->effective_window_size + r.cycle_samples);
			first[s] = (long long)n_samples - (long long)offset2;
		}
		SGZS_CUDA(cudaMemcpyAsync(dFirst, first.data(), n_streams * sizeof(long long), cudaMemcpyHostToDevice, c->stream));
		SGZS_CUDA(cudaMemcpyAsync(dRadians, radians.data(), n_streams * sizeof(double), cudaMemcpyHostToDevice, c->stream));
		GoertzelArgs ga{};
		ga.audio = d; ga.layout = layout; ga.n = n_samples; ga.stride = stream_stride; ga.channel = prm->channel; ga.firstSample = dFirst; ga.radians = dRadians; ga.z = dZ;
		k_scope_goertzel<<<n_streams, 256, 0, c->stream>>>(ga);
		SGZS_CUDA(cudaGetLastError());
		std::vector<double2> z(n_streams);
		SGZS_CUDA(cudaMemcpyAsync(z.data(), dZ, n_streams * sizeof(double2), cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		for (uint32_t s = 0; s < n_streams; ++s)
		{
			// :283-306
			sgz_scope_period_result& r = results[s];
			const double rotation = -sampleDifference[s] * radians[s];
			const double cr = std::cos(rotation), ci = -std::sin(rotation);
			const double zr = z[s].x * cr - z[s].y * ci, zi = z[s].x * ci + z[s].y * cr;
			double phase = tau - std::atan2(zi, zr);
			phase += r.record.offset * tau;
			phase -= 1.5707963267948966192313216916398;
			phase += tau * prm->phase_offset_degrees / 360;
			phase = std::fmod(phase, tau);
			while (phase < 0) phase += tau;
			r.phase = phase;
			r.sample_offset = (phase / tau) * prm->sample_rate / r.fundamental - 1;
		}
		return SGZ_OK;
	}

	// ---- oscilloscope colouring and envelopes --------------------------------------------------
	static void designSections(double fNorm, ColourSection& lp, ColourSection& hp)
	{
		// 2nd-order Butterworth pair through the bilinear transform; two passes of each make the 4th-order
		// Linkwitz-Riley crossover (cpl decisions)
		const double w0 = 6.283185307179586476925286766559 * fNorm, c = std::cos(w0), alpha = std::sin(w0) / 1.4142135623730950488016887242097;
		const double norm = 1.0 + alpha;
		lp.b0 = float((1 - c) * 0.5 / norm); lp.b1 = float((1 - c) / norm); lp.b2 = lp.b0;
		hp.b0 = float((1 + c) * 0.5 / norm); hp.b1 = float(-(1 + c) / norm); hp.b2 = hp.b0;
		lp.a1 = hp.a1 = float(-2 * c / norm);
		lp.a2 = hp.a2 = float((1 - alpha) / norm);
	}

	int sgz_scope_colour(sgz_context* c, const sgz_scope_colour_config* cfg, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, sgz_scope_colour_state* states, uint8_t* rgba, int rgba_mem, double* envelope_gain)
	{
		int rc;
		if (!cfg || cfg->abi_version != SGZ_ABI_VERSION || !states || !rgba) return fail(SGZ_EINVAL, "bad colour arguments");
		if (cfg->channel_mode < SGZ_OSC_LEFT || cfg->channel_mode > SGZ_OSC_MIDSIDE) return fail(SGZ_EINVAL, "bad channel mode");
		if (cfg->envelope_mode < SGZ_ENVELOPE_NONE || cfg->envelope_mode > SGZ_ENVELOPE_PEAK) return fail(SGZ_EINVAL, "bad envelope mode");
		if (!(cfg->sample_rate > 0)) return fail(SGZ_EINVAL, "sample_rate must be positive");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		if (n_samples == 0) return SGZ_OK;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		const size_t stateBytes = size_t(n_streams) * 2 * sizeof(sgz_scope_colour_state);
		void* dStates = nullptr; if ((rc = ensure(c, 2, stateBytes, &dStates)) != SGZ_OK) return rc;
		SGZS_CUDA(cudaMemcpyAsync(dStates, states, stateBytes, cudaMemcpyHostToDevice, c->stream));
		const size_t bytes = size_t(n_streams) * 4 * n_samples * 4;
		uint8_t* dOut = rgba;
		if (rgba_mem == SGZ_MEM_HOST) { void* p = nullptr; if ((rc = ensure(c, 1, bytes, &p)) != SGZ_OK) return rc; dOut = static_cast<uint8_t*>(p); }
		if (reinterpret_cast<uintptr_t>(dOut) & 3) return fail(SGZ_EINVAL, "rgba must be 4-byte aligned");
		ColourArgs a{};
		a.audio = d; a.layout = layout; a.n = n_samples; a.stride = stream_stride; a.nStreams = n_streams;
		designSections(double(static_cast<float>(300.0 / cfg->sample_rate)), a.lp[0], a.hp[0]);   // tuneCrossOver(300, 3000, sampleRate), ChannelData.h:163-166
		designSections(double(static_cast<float>(3000.0 / cfg->sample_rate)), a.lp[1], a.hp[1]);
		a.pole = static_cast<float>(std::exp(-1.0 / (cfg->colour_smoothing_ms * 0.001 * cfg->sample_rate)));
		a.envelopeCoeff = static_cast<float>(std::exp(-1.0 / (cfg->envelope_window * cfg->sample_rate))); // :448
		a.blend = 1 - static_cast<float>(cfg->colouring_blend);                                            // :503
		for (int i = 0; i < 3; ++i) for (int k = 0; k < 3; ++k) a.colours[i][k] = cfg->colours[i][k];
		for (int k = 0; k < 4; ++k) { a.keyL[k] = float(cfg->key_left[k]); a.keyR[k] = float(cfg->key_right[k]); }
		a.channelMode = cfg->channel_mode; a.envelopeMode = cfg->envelope_mode;
		a.states = static_cast<sgz_scope_colour_state*>(dStates); a.rgba = dOut;
		// few pairs: a warp per pair (recursions on two lanes, colouring by all 32); thousands of pairs: the recursions
		// alone fill the GPU, so pack 16 pairs per warp (measured crossover between 256 and 4096 pairs)
		if (n_streams >= kColourPackedFrom) k_scope_colour_packed<<<(2 * n_streams + 63) / 64, 64, 0, c->stream>>>(a);
		else k_scope_colour<<<(n_streams + kColourWarps - 1) / kColourWarps, 32 * kColourWarps, 0, c->stream>>>(a);
		SGZS_CUDA(cudaGetLastError());
		SGZS_CUDA(cudaMemcpyAsync(states, dStates, stateBytes, cudaMemcpyDeviceToHost, c->stream));
		if (rgba_mem == SGZ_MEM_HOST) SGZS_CUDA(cudaMemcpyAsync(rgba, dOut, bytes, cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		if (envelope_gain)
			for (uint32_t s = 0; s < n_streams; ++s)
			{
				// :682-699: 1 / the larger of the two channels' RMS
				double g = 0;
				if (cfg->envelope_mode == SGZ_ENVELOPE_RMS)
					g = 1.0 / std::max(std::sqrt(states[2 * s].envelope), std::sqrt(states[2 * s + 1].envelope));
				envelope_gain[s] = g;
			}
		return SGZ_OK;
	}

	int sgz_scope_peak_envelope(sgz_context* c, int channel_mode, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples,
		uint64_t stream_stride, double coeff, float* envelopes, double* auto_gain)
	{
		int rc;
		if (!envelopes || !auto_gain) return fail(SGZ_EINVAL, "null argument");
		if (channel_mode < SGZ_OSC_LEFT || channel_mode > SGZ_OSC_MIDSIDE) return fail(SGZ_EINVAL, "bad channel mode");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		void* dMax = nullptr; if ((rc = ensure(c, 2, size_t(n_streams) * sizeof(float2), &dMax)) != SGZ_OK) return rc;
		const uint64_t stop = n_samples - (n_samples & 7); // whole 8-float vectors only (:738-740)
		k_scope_peak_max<<<n_streams, 256, 0, c->stream>>>(d, layout, n_samples, stream_stride, stop, channel_mode, static_cast<float2*>(dMax));
		SGZS_CUDA(cudaGetLastError());
		std::vector<float2> hi(n_streams);
		SGZS_CUDA(cudaMemcpyAsync(hi.data(), dMax, size_t(n_streams) * sizeof(float2), cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		for (uint32_t s = 0; s < n_streams; ++s)
		{
			float* env = envelopes + 2 * size_t(s);
			if (channel_mode == SGZ_OSC_SEPARATE)
			{
				// float arithmetic here (std::max<float>), and the second channel inherits the first one's running maximum
				const float first = hi[s].x, second = std::max(hi[s].x, hi[s].y);
				env[0] = std::max<float>(float(env[0] * coeff), first * first);
				env[1] = std::max<float>(float(env[1] * coeff), second * second);
			}
			else
			{
				const double a = hi[s].x, b = hi[s].y;
				env[0] = static_cast<float>(std::max(env[0] * coeff, a * a));
				env[1] = static_cast<float>(std::max(env[1] * coeff, b * b));
			}
			auto_gain[s] = 1.0 / std::max(std::sqrt(env[0]), std::sqrt(env[1])); // :877-885
		}
		return SGZ_OK;
	}
	// VectorScope::runPeakFilter (VectorscopeRendering.cpp:826-889): largest |L| and |R| of the window (whole groups of 8
	// samples only, the AVX loop of :857-868), squared and folded into processor->filters.envelope[0 / 1] with
	// coeff = envelopeCoeff^(numSamples * openGLDeltaTime) (:840-842), then envelopeGain = 1 / max(sqrt, sqrt), kept only
	// when it is a normal number (:882-888).  The window maximum runs on the GPU (k_scope_peak_max in its two-channel
	// form); the fold is a handful of scalar operations per stream on the host, in the reference's types (float envelope
	// states, double arithmetic, float square roots).
	int sgz_vscope_peak_envelope(sgz_context* c, const float* audio, int mem, int layout, uint32_t n_streams, uint64_t n_samples, uint64_t stream_stride,
		float envelope_coeff, double delta_time, sgz_vscope_filters* filters, double* envelope_gain)
	{
		int rc;
		if (!filters || !envelope_gain) return fail(SGZ_EINVAL, "null argument");
		if ((rc = validateAudio(audio, layout, n_streams, n_samples, stream_stride)) != SGZ_OK) return rc;
		SGZS_GUARD(c);
		const float* d = nullptr;
		if ((rc = stageIn(c, audio, mem, n_streams, n_samples, stream_stride, &d)) != SGZ_OK) return rc;
		void* dMax = nullptr; if ((rc = ensure(c, 2, size_t(n_streams) * sizeof(float2), &dMax)) != SGZ_OK) return rc;
		const uint64_t stop = n_samples - (n_samples & 7); // whole 8-float vectors only (:860); "TODO: remainder?" (:870)
		k_scope_peak_max<<<n_streams, 256, 0, c->stream>>>(d, layout, n_samples, stream_stride, stop, SGZ_OSC_SEPARATE, static_cast<float2*>(dMax));
		SGZS_CUDA(cudaGetLastError());
		std::vector<float2> hi(n_streams);
		SGZS_CUDA(cudaMemcpyAsync(hi.data(), dMax, size_t(n_streams) * sizeof(float2), cudaMemcpyDeviceToHost, c->stream));
		SGZS_CUDA(cudaStreamSynchronize(c->stream));
		const double power = double(n_samples) * delta_time;
		const double coeff = std::pow(envelope_coeff, power); // std::pow(float, double) -> double
		for (uint32_t s = 0; s < n_streams; ++s)
		{
			float* env = filters[s].envelope;
			const double highestLeft = hi[s].x, highestRight = hi[s].y;
			env[0] = static_cast<float>(std::max(env[0] * coeff, highestLeft * highestLeft));
			env[1] = static_cast<float>(std::max(env[1] * coeff, highestRight * highestRight));
			const double currentEnvelope = 1.0 / std::max(std::sqrt(env[0]), std::sqrt(env[1]));
			if (std::isnormal(currentEnvelope)) envelope_gain[s] = currentEnvelope;
		}
		return SGZ_OK;
	}
	// TriggeringProcessor::update + ::processMutating (StreamPreprocessing.h:54-75, :78-206) for every stream of a batch.
	// The peak queue is the caller's row [head, count); events stand in for the two side effects (back-buffer append,
	// front-buffer swap).  The decisions mix std::uint64_t clocks with the double halfSize exactly as the reference
	// writes them (comparisons and sums in double, truncation where it assigns to an integer).
	namespace
	{
		struct WindowRun
		{
			sgz_scope_window_state st;
			const uint64_t* queue; uint32_t head, count; // peaks.front() == queue[head]
			uint64_t numSamples;
			sgz_scope_window_event* events; uint32_t nEvents, maxEvents;
			bool overflow;
			double ceilingSize;

			bool hasPeaks() const { return head < count; }
			void emit(uint32_t kind, uint64_t origin, uint64_t size)
			{
				if (nEvents < maxEvents) events[nEvents] = sgz_scope_window_event{ origin, size, st.buffered_samples, kind, 0u };
				else overflow = true;
				++nEvents;
			}
			void processIntoBackBuffer(uint64_t samples) // the lambda at :89-103
			{
				emit(SGZ_WINDOW_PROCESS, st.steady_clock, samples);
				numSamples -= samples;
				const uint64_t oldSamples = st.buffered_samples;
				st.steady_clock += samples;
				st.buffered_samples += samples;
				st.buffered_samples = std::min<uint64_t>(st.buffered_samples, static_cast<uint64_t>(ceilingSize + 1));
				st.front_origin += (oldSamples + samples) - st.buffered_samples;
			}
		};
	}

	int sgz_scope_trigger_windows(uint32_t n_streams, sgz_scope_window_state* states, uint64_t* peaks, uint32_t* peak_counts, uint32_t max_peaks,
		uint64_t steady_clock, uint64_t n_samples, sgz_scope_window_event* events, uint32_t* event_counts, uint32_t max_events)
	{
		if (!states || !peak_counts || !event_counts || (!peaks && max_peaks) || (!events && max_events)) return fail(SGZ_EINVAL, "null argument");
		for (uint32_t s = 0; s < n_streams; ++s) if (peak_counts[s] > max_peaks) return fail(SGZ_EINVAL, "stream %u: peak count %u exceeds max_peaks %u", s, peak_counts[s], max_peaks);
		for (uint32_t s = 0; s < n_streams; ++s)
		{
			WindowRun r{};
			r.st = states[s];
			r.queue = peaks + size_t(s) * max_peaks; r.head = 0; r.count = peak_counts[s];
			r.numSamples = n_samples;
			r.events = events + size_t(s) * max_events; r.maxEvents = max_events;
			sgz_scope_window_state& st = r.st;

			// update(currentSteadyClock), :54-75
			st.steady_clock = steady_clock;
			if (st.window_changed)
			{
				st.window_changed = 0;
				if (st.is_working_on_peak && r.hasPeaks()) ++r.head;
				while (r.hasPeaks() && r.queue[r.head] < steady_clock) ++r.head;
				st.buffered_samples = st.current_peak = st.old_peak = 0;
				st.front_origin = steady_clock;
				st.is_working_on_peak = 0;
			}

			// processMutating, :78-206
			if (st.front_origin + st.buffered_samples < st.steady_clock)
			{
				st.front_origin = st.steady_clock;
				st.buffered_samples = 0;
			}
			r.ceilingSize = std::ceil(st.window_size);
			const double halfSize = r.ceilingSize / 2;
			if (r.ceilingSize == 0 && r.hasPeaks()) r.head = r.count;
			while (r.numSamples != 0)
			{
				if (!r.hasPeaks())
				{
					r.processIntoBackBuffer(r.numSamples);
					break;
				}
				else if (!st.is_working_on_peak)
				{
					st.is_working_on_peak = 1;
					const uint64_t nextPeak = r.queue[r.head];
					if (nextPeak >= st.steady_clock)
					{
						// select closest peak that has a full buffer
						const uint64_t deltaToPeak = nextPeak - st.steady_clock;
						r.processIntoBackBuffer(std::min<uint64_t>(r.numSamples, static_cast<uint64_t>(double(deltaToPeak) + halfSize)));
					}
					st.current_peak = nextPeak;
				}
				const uint64_t sincePeak = st.current_peak - st.old_peak;
				const bool isPeakOutsideOfWindow = !(double(sincePeak) < halfSize);
				const uint64_t windowEnd = isPeakOutsideOfWindow ? st.front_origin + st.buffered_samples : static_cast<uint64_t>(double(st.old_peak) + halfSize);
				const uint64_t peakWindowEnd = static_cast<uint64_t>(double((isPeakOutsideOfWindow ? 1 : 0) + st.current_peak) + halfSize);
				const uint64_t missingBufferSamples = peakWindowEnd - std::min(peakWindowEnd, windowEnd);
				const uint64_t neededPreSamples = std::min<uint64_t>(static_cast<uint64_t>(halfSize), static_cast<uint64_t>(std::max<double>(double(sincePeak), halfSize) - halfSize));
				bool readyForBufferSwap;
				if (isPeakOutsideOfWindow)
				{
					const uint64_t n = std::min<uint64_t>(r.numSamples, missingBufferSamples);
					if (n > 0) r.processIntoBackBuffer(n);
					readyForBufferSwap = missingBufferSamples == n;
				}
				else if (st.buffered_samples >= missingBufferSamples) readyForBufferSwap = true;
				else
				{
					const uint64_t numRemaining = missingBufferSamples - st.buffered_samples;
					const uint64_t n = std::min<uint64_t>(r.numSamples, numRemaining);
					if (n > 0) r.processIntoBackBuffer(n);
					readyForBufferSwap = numRemaining == n;
				}
				if (readyForBufferSwap)
				{
					const double amount = (isPeakOutsideOfWindow ? halfSize : double(missingBufferSamples)) + double(neededPreSamples);
					const uint64_t cappedSize = std::min<size_t>(st.buffered_samples, static_cast<size_t>(std::ceil(amount + 1)));
					r.emit(SGZ_WINDOW_SWAP, st.front_origin, cappedSize);
					st.buffered_samples -= std::min<uint64_t>(st.buffered_samples, cappedSize);
					st.front_origin += cappedSize;
					st.old_peak = st.current_peak;
					st.is_working_on_peak = 0;
					++r.head;
				}
			}
			if (r.overflow) return fail(SGZ_EINVAL, "stream %u needs %u window events, max_events is %u (nothing was written back for it)", s, r.nEvents, max_events);
			states[s] = st;
			event_counts[s] = r.nEvents;
			uint64_t* row = peaks + size_t(s) * max_peaks;
			for (uint32_t k = r.head; k < r.count; ++k) row[k - r.head] = row[k];
			peak_counts[s] = r.count - r.head;
		}
		return SGZ_OK;
	}
}
