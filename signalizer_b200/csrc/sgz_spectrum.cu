// sgz_spectrum.cu — sm_100a kernels and the C ABI (include/sgz.h) of the Spectrum transform path.
//
// Reference path (relative to the reference checkout):
//   Source/Spectrum/TransformDSP.inl   prepareTransform :39-484, doTransform :487-502,
//                                      mapToLinearSpace :506-1135, mapAndTransformDFTFilters :1299-1435
//   Source/Spectrum/TransformConstant.h, TransformPair.h
//
// Two implementations of the same arithmetic live here:
//   * the staged kernels (k_prepare, k_fft_*, k_map, k_post) mirror the reference's four calls one to
//     one, for every SpectrumChannels x BinInterpolation combination and every power-of-two size,
//     keeping the intermediate buffers (audioMemory / workingMemory of TransformPair.h:238-243) in HBM;
//   * k_fused runs a whole frame (all four calls) per CTA without touching HBM in between, for the
//     mono and Separate/MidSide configurations at N = 4096 / 8192, looping over the frames of one
//     stream so the peak-hold state stays on chip.
// There is no CPU fallback anywhere in this file.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <type_traits>
#include <atomic>
#include <mutex>
#include <complex>
#include <vector>

#include "../../include/sgz.h"
#include "sgz_device.cuh"
#include "sgz_tables.hpp"

namespace sgz
{
	// first-pass radix of the fused kernel k_fused8<K1> for transform size N (0 = no fused kernel at this size)
	inline int fusedK1For(uint32_t N) { return N == 16384 ? 16 : N == 8192 ? 8 : N == 4096 ? 4 : N == 2048 ? 2 : N == 1024 ? 1 : 0; }
	// the fused kernel keeps 28 bytes per display point next to the N * 8-byte exchange buffer; one CTA must fit the
	// 227 KB an SM can give (the resident CTAs per SM then follow from the size: 2 at N = 8192 up to 2048 points, 1 beyond)
	inline bool fusedSharedMemoryFits(uint32_t N, uint32_t P) { return size_t(N) * 8 + 2048 + 80 + size_t(P) * 28 <= size_t(227) * 1024; }
	constexpr uint32_t kFrameParallelMinRun = 4;   // with fewer streams than resident CTAs a stream is cut into parts of at least this many frames

	// ------------------------------------------------------------------------------------------
	// error plumbing
	// ------------------------------------------------------------------------------------------
	static thread_local std::string g_lastError;

	// records the message behind sgz_last_error(); shared with the other .cu files of the library
	int fail(int code, const char* fmt, ...)
	{
		char buf[512];
		va_list ap;
		va_start(ap, fmt);
		vsnprintf(buf, sizeof(buf), fmt, ap);
		va_end(ap);
		g_lastError = buf;
		return code;
	}

#define SGZ_CUDA(expr)                                                                                      \
	do {                                                                                                    \
		cudaError_t e__ = (expr);                                                                           \
		if (e__ != cudaSuccess)                                                                             \
			return fail(SGZ_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
	} while (0)

	// ------------------------------------------------------------------------------------------
	// device-side view of the tables (all pointers are device memory owned by the handle)
	// ------------------------------------------------------------------------------------------
	struct DeviceTables
	{
		const float* window;     // [N]
		const float* slope;      // [P]
		const int4* binmap;      // [P] {mode, first, count, fallback}
		const float* posL;       // [P]
		const float* posR;       // [P]
		const int* baseL;        // [P]
		const int* baseR;        // [P]
		const float* wL;         // [P][10]
		const float* wR;         // [P][10]
		const float4* wL4;       // [3][P] the same weights, four taps per word, point-major within a word row (fused kernel)
		const float4* wR4;
		const int* phaseNormPos; // [P]
		const float2* twN;       // [N/2] W_N^k (generic FFT)
		const float2* tw3;       // [M]      W_N^k
		const float2* tw8;       // k_fused8: [8][8] W_64^(ab) | [8][64] W_M^(v k1) | [8][64] W_512^(v k2)
		const float2* resCoeff;  // [V][P] pole per vector and filter
		const float* resGain;    // [P]
		const float* resTaps;    // [V]
		uint32_t resVectors, resChannels;
		int phaseNormFinal;
		uint32_t N, W, P, M, linLimit;
		uint32_t workStride; // float2 entries of workingMemory per stream
		uint32_t safeFrom;   // fused kernel: interpolated points x >= safeFrom need no tap bounds checks
		int conf, binpol;
		float invSize, deltaYRecip, minFracRecip, lowerClip;
		float deltaYLn2; // deltaYRecip * ln 2, for the log2-based display scale of the fused kernel
		float pole0, pole1, phasePole0, phasePole1;
		// cosine-sum window spanning the whole transform (W == N, periodic): w(n) = winPoly(cos(2 pi n / N)), powers 0..4;
		// winStep = (cos, sin)(2 pi / N)
		float winPoly[5];
		float2 winStep;
		int winPolyOk;
	};

	__device__ __forceinline__ bool isMonoConf(int conf) { return conf <= SGZ_CH_SIDE; }

	// channel matrix of prepareTransform (TransformDSP.inl:59-217), multiplication order as written
	__device__ __forceinline__ float2 channelMatrix(int conf, float l, float r, float w)
	{
		switch (conf)
		{
		case SGZ_CH_LEFT: return make_float2(l * w, 0.f);
		case SGZ_CH_RIGHT: return make_float2(r * w, 0.f);
		case SGZ_CH_MERGE: return make_float2((l + r) * w * 0.5f, 0.f);
		case SGZ_CH_SIDE: return make_float2((l - r) * w * 0.5f, 0.f);
		case SGZ_CH_MIDSIDE: return make_float2((l + r) * w * 0.5f, (l - r) * w * 0.5f);
		default: return make_float2(l * w, r * w);
		}
	}

	// ------------------------------------------------------------------------------------------
	// staged kernel 1: prepareTransform
	// ------------------------------------------------------------------------------------------
	struct PrepareArgs
	{
		const float* hist; uint64_t nHist; uint64_t histStride;     // history, per stream
		const float* prelim; uint64_t nPrelim; uint64_t prelimStride; // preliminary audio (may be null)
		int layout;
		uint32_t histSkip;   // index of the first history sample used
		uint32_t histCount;  // number of window taps filled from history
		uint32_t prelimCount; // number of taps filled from preliminary audio ("stop")
	};

	__device__ __forceinline__ void fetchLR(const float* base, int layout, uint64_t n, uint64_t i, float& l, float& r)
	{
		if (layout == SGZ_LAYOUT_INTERLEAVED) { const float2 v = reinterpret_cast<const float2*>(base)[i]; l = v.x; r = v.y; }
		else { l = base[i]; r = base[n + i]; }
	}

	__global__ void k_prepare(DeviceTables t, PrepareArgs a, float2* __restrict__ csf)
	{
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t s = blockIdx.y;
		if (i > t.N) return;
		float2 out = make_float2(0.f, 0.f);
		if (i < a.histCount)
		{
			float l, r;
			fetchLR(a.hist + s * a.histStride, a.layout, a.nHist, uint64_t(a.histSkip) + i, l, r);
			out = channelMatrix(t.conf, l, r, t.window[i]);
		}
		else if (i < a.histCount + a.prelimCount)
		{
			float l, r;
			fetchLR(a.prelim + s * a.prelimStride, a.layout, a.nPrelim, uint64_t(i - a.histCount), l, r);
			out = channelMatrix(t.conf, l, r, t.window[i]);
		}
		csf[uint64_t(s) * (t.N + 1) + i] = out; // zero-pad, including the extra element N
	}

	// audio history ring: sample `skip + j` of the block goes to ring position (head + skip + j) mod C and to
	// its mirror at + C
	__global__ void k_history_append(float2* __restrict__ ring, uint64_t C, uint64_t head, const float* __restrict__ block, int layout, uint64_t nBlock,
		uint64_t stride, uint64_t skip)
	{
		const uint64_t j = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
		if (skip + j >= nBlock) return;
		const uint32_t s = blockIdx.y;
		float l, r;
		fetchLR(block + s * stride, layout, nBlock, skip + j, l, r);
		const uint64_t pos = (head + skip + j) % C;
		float2* dst = ring + uint64_t(s) * 2 * C;
		dst[pos] = make_float2(l, r);
		dst[pos + C] = make_float2(l, r);
	}

	// ------------------------------------------------------------------------------------------
	// frequency tracker: per stream, first maximum of key(i) over [lo, hiEx), then the reference's
	// "continuously rising" walk when the maximum sits on a border of the window
	// (SpectrumRendering.cpp:320-350, 391-421).  key = .leftMagnitude of a results graph, or |z|^2 of
	// the raw FFT.  One warp per stream; the handful of scalars the caller derives from the peak
	// (parabolic fit, dB conversion, smoothing) are finished on the host from the record written here.
	// ------------------------------------------------------------------------------------------
	struct TrackBounds { long long lo, hiEx, rightBorder, end; };
	struct TrackRecord { long long peak; float2 prev, at, next; };

	template<bool RAW>
	__device__ __forceinline__ float trackKey(const float2* d, long long i)
	{
		const float2 v = d[i];
		if (RAW) return __fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y));
		return v.x;
	}

	template<bool RAW>
	__global__ void k_track_peaks(const float2* __restrict__ data, uint64_t stride, const TrackBounds* __restrict__ bounds, TrackRecord* __restrict__ out, uint32_t nStreams)
	{
		const uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
		const int lane = threadIdx.x & 31;
		if (s >= nStreams) return;
		const float2* d = data + uint64_t(s) * stride;
		const TrackBounds b = bounds[s];
		float best = -INFINITY;
		long long at = 0x7fffffffffffffffll;
		for (long long i = b.lo + lane; i < b.hiEx; i += 32)
		{
			const float v = trackKey<RAW>(d, i);
			if (at == 0x7fffffffffffffffll || best < v) { best = v; at = i; } // strict: the earliest of equal maxima stays
		}
#pragma unroll
		for (int o = 16; o > 0; o >>= 1)
		{
			const float ov = __shfl_xor_sync(0xffffffffu, best, o);
			const long long oa = __shfl_xor_sync(0xffffffffu, at, o);
			const bool mine = at != 0x7fffffffffffffffll, theirs = oa != 0x7fffffffffffffffll;
			if (theirs && (!mine || best < ov || (best == ov && oa < at))) { best = ov; at = oa; }
		}
		if (lane != 0) return;
		long long peak = at;
		if (peak == b.lo && b.lo != 0)
		{
			while (true)
			{
				const long long next = peak - 1;
				if (next == 0) break;
				if (trackKey<RAW>(d, next) < trackKey<RAW>(d, peak)) break;
				peak = next;
			}
		}
		else if (peak == b.rightBorder)
		{
			while (true)
			{
				const long long next = peak + 1;
				if (next == b.end) break;
				if (trackKey<RAW>(d, next) < trackKey<RAW>(d, peak)) break;
				peak = next;
			}
		}
		TrackRecord r;
		r.peak = peak;
		r.at = d[peak];
		r.prev = d[peak == 0 ? 0 : peak - 1];
		// RAW: the reference reads source[peak + 1] up to index N, which audioMemory carries; results graphs end at P - 1
		r.next = d[RAW ? (peak == b.end ? peak : peak + 1) : (peak + 1 < b.end ? peak + 1 : peak)];
		out[s] = r;
	}

	// ------------------------------------------------------------------------------------------
	// staged kernel 2: doTransform — generic power-of-two forward FFT, natural order, in place
	// ------------------------------------------------------------------------------------------
	// N <= 16384 (other than the sizes k_fft8 takes): one CTA per stream, decimation in time in shared memory on the
	// bit-reversed input, two radix-2 stages per sweep (one barrier and one round trip through shared memory for both);
	// an odd log2 N starts with a single radix-2 stage
	__global__ void __launch_bounds__(256) k_fft_smem(DeviceTables t, float2* __restrict__ csf)
	{
		extern __shared__ float2 sm[];
		const uint32_t N = t.N;
		float2* x = csf + uint64_t(blockIdx.x) * (N + 1);
		const int bits = 31 - __clz(N);
		for (uint32_t i = threadIdx.x; i < N; i += blockDim.x)
			sm[__brev(i) >> (32 - bits)] = x[i];
		__syncthreads();
		uint32_t half = 1;
		if (bits & 1)
		{
			for (uint32_t b = threadIdx.x; b < (N >> 1); b += blockDim.x)
			{
				const float2 u = sm[2 * b], v = sm[2 * b + 1]; // W_2^0 = 1
				sm[2 * b] = cadd(u, v);
				sm[2 * b + 1] = csub(u, v);
			}
			__syncthreads();
			half = 2;
		}
		for (; half < N; half <<= 2)
		{
			// stages of size 2 half and 4 half on the four elements base + {0, 1, 2, 3} half
			const uint32_t step1 = (N >> 1) / half, step2 = step1 >> 1; // W_(2 half)^j = W_N^(j step1), W_(4 half)^j = W_N^(j step2)
			for (uint32_t b = threadIdx.x; b < (N >> 2); b += blockDim.x)
			{
				const uint32_t j = b & (half - 1);
				const uint32_t base = ((b - j) << 2) + j;
				const float2 w1 = t.twN[j * step1], w2 = t.twN[j * step2];
				const float2 a0 = sm[base], a1 = cmul(sm[base + half], w1), a2 = sm[base + 2 * half], a3 = cmul(sm[base + 3 * half], w1);
				const float2 u0 = cadd(a0, a1), u1 = csub(a0, a1);
				const float2 u2 = cmul(cadd(a2, a3), w2), u3 = cmul(csub(a2, a3), w2);
				sm[base] = cadd(u0, u2);
				sm[base + 2 * half] = csub(u0, u2);
				// the odd outputs take W_(4 half)^(j + half) = -i W_(4 half)^j: -i (x + i y) = y - i x
				const float2 r3 = make_float2(u3.y, -u3.x);
				sm[base + half] = cadd(u1, r3);
				sm[base + 3 * half] = csub(u1, r3);
			}
			__syncthreads();
		}
		for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) x[i] = sm[i];
	}

	// N > 16384: Stockham radix-2 passes through HBM (ping-pong with the work buffer)
	__global__ void k_fft_global_pass(const float2* __restrict__ tw, const float2* __restrict__ src, float2* __restrict__ dst,
		uint32_t N, uint32_t n, uint32_t s, uint64_t srcStride, uint64_t dstStride)
	{
		// one thread per butterfly: p in [0, n/2), q in [0, s): b = p * s + q
		const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
		if (b >= (N >> 1)) return;
		const float2* x = src + uint64_t(blockIdx.y) * srcStride;
		float2* y = dst + uint64_t(blockIdx.y) * dstStride;
		const uint32_t q = b % s, p = b / s;
		const uint32_t n1 = n >> 1;
		const float2 w = tw[p * s]; // W_n^p = W_N^(p s)
		const float2 u = x[q + s * p], v = x[q + s * (p + n1)];
		y[q + s * (2 * p)] = cadd(u, v);
		y[q + s * (2 * p + 1)] = cmul(csub(u, v), w);
	}

	__global__ void k_copy_strided(const float2* __restrict__ src, float2* __restrict__ dst, uint32_t N, uint64_t srcStride, uint64_t dstStride)
	{
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < N) dst[uint64_t(blockIdx.y) * dstStride + i] = src[uint64_t(blockIdx.y) * srcStride + i];
	}

	// ------------------------------------------------------------------------------------------
	// staged kernel 3: mapToLinearSpace (FFT branch), one CTA per stream, csf modified in place like
	// the reference does.
	// ------------------------------------------------------------------------------------------
	__device__ __forceinline__ float2 tapOr0(const float2* csf, int i, int N) { return (i >= 0 && i <= N) ? csf[i] : make_float2(0.f, 0.f); }

	__device__ __forceinline__ float2 linearAt(const float2* csf, float pos, int N)
	{
		const float fl = floorf(pos);
		const int i = int(fl);
		const float frac = pos - fl;
		const float2 a = tapOr0(csf, i, N), b = tapOr0(csf, i + 1, N);
		return make_float2(a.x * (1.f - frac) + b.x * frac, a.y * (1.f - frac) + b.y * frac);
	}

	__device__ __forceinline__ float2 lanczosAt(const float2* csf, int base, const float* w, int N)
	{
		float2 acc = make_float2(0.f, 0.f);
#pragma unroll
		for (int k = 0; k < kLanczosTaps; ++k)
		{
			const float2 v = tapOr0(csf, base + k, N);
			acc.x += v.x * w[k]; acc.y += v.y * w[k];
		}
		return acc;
	}

	// Phase configuration: value of csf[i] as the reference's magnitude pass sees it when `np` low
	// indices (and their mirrors) have been replaced by their magnitude (TransformDSP.inl:715-720)
	__device__ __forceinline__ float2 phaseView(const float2* csf, int i, int N, int np)
	{
		if (i < 0 || i > N) return make_float2(0.f, 0.f);
		const float2 v = csf[i];
		return (i < np || i > N - np) ? make_float2(cabs(v), 0.f) : v;
	}

	__global__ void __launch_bounds__(256) k_map(DeviceTables t, float2* __restrict__ csfAll, float2* __restrict__ workAll)
	{
		const int N = int(t.N), M = int(t.M), P = int(t.P);
		float2* csf = csfAll + uint64_t(blockIdx.x) * (N + 1);
		const int conf = t.conf;
		const bool mono = isMonoConf(conf);
		const bool phase = conf == SGZ_CH_PHASE;
		const bool cplx = conf == SGZ_CH_COMPLEX;
		const int chans = mono || cplx ? 1 : 2;
		float2* work = workAll + uint64_t(blockIdx.x) * t.workStride;

		// ---- phase A: in-place normalisation ----
		if (mono)
		{
			// :553-560  csf[0] *= .5; csf[N/2] *= .5; csf[i] = |csf[i]| for i < numBins
			for (int i = threadIdx.x; i < M; i += blockDim.x)
			{
				float2 v = csf[i];
				if (i == 0) { v.x *= 0.5f; v.y *= 0.5f; }
				csf[i] = make_float2(cabs(v), 0.f);
			}
			if (threadIdx.x == 0) { float2 v = csf[M]; csf[M] = make_float2(v.x * 0.5f, v.y * 0.5f); }
		}
		else if (cplx)
		{
			// :993-1002
			for (int i = threadIdx.x; i < N; i += blockDim.x)
			{
				float2 v = csf[i];
				if (i == 0) csf[0] = make_float2(v.x * 0.5f, v.y * 0.5f);
				else csf[i] = make_float2(cabs(v), 0.f);
			}
		}
		else
		{
			// two-for-one split (cpl::dsp::separateTransformsIPL) + :861-869 / :649-652
			for (int k = threadIdx.x; k < M; k += blockDim.x)
			{
				if (k == 0)
				{
					const float2 z0 = csf[0], zm = csf[M];
					csf[N] = make_float2(z0.y * 0.5f, 0.f);
					csf[0] = make_float2(z0.x * 0.5f, 0.f);
					const float2 h = make_float2(zm.x * 0.5f, zm.y * 0.5f);
					csf[M] = phase ? h : make_float2(cabs(h), 0.f);
					continue;
				}
				const float2 zk = csf[k], zn = csf[N - k];
				float2 A = make_float2((zk.x + zn.x) * 0.5f, (zk.y - zn.y) * 0.5f);
				const float2 D = make_float2((zk.x - zn.x) * 0.5f, (zk.y + zn.y) * 0.5f);
				const float2 B = make_float2(D.y, -D.x);
				if (k == M - 1) { A.x *= 0.5f; A.y *= 0.5f; }
				if (phase) { csf[k] = A; csf[N - k] = B; }
				else { csf[k] = make_float2(cabs(A), 0.f); csf[N - k] = make_float2(cabs(B), 0.f); }
			}
		}
		__syncthreads();

		// ---- phase B: display points ----
		const float invSize = t.invSize;
		for (int x = threadIdx.x; x < P; x += blockDim.x)
		{
			const int4 bm = t.binmap[x];
			if (phase)
			{
				float mid, cancelOut;
				if (bm.x == 0)
				{
					float2 iL, iR, mL, mR;
					if (t.binpol == SGZ_BIN_NONE)
					{
						iL = csf[bm.y]; iR = csf[N - bm.y]; mL = iL; mR = iR;
					}
					else if (t.binpol == SGZ_BIN_LINEAR)
					{
						iL = linearAt(csf, t.posL[x], N); iR = linearAt(csf, t.posR[x], N);
						const int np = t.phaseNormPos[x];
						{
							const float pos = t.posL[x]; const float fl = floorf(pos); const int i = int(fl); const float fr = pos - fl;
							const float2 a = phaseView(csf, i, N, np), b = phaseView(csf, i + 1, N, np);
							mL = make_float2(a.x * (1.f - fr) + b.x * fr, a.y * (1.f - fr) + b.y * fr);
						}
						{
							const float pos = t.posR[x]; const float fl = floorf(pos); const int i = int(fl); const float fr = pos - fl;
							const float2 a = phaseView(csf, i, N, np), b = phaseView(csf, i + 1, N, np);
							mR = make_float2(a.x * (1.f - fr) + b.x * fr, a.y * (1.f - fr) + b.y * fr);
						}
					}
					else
					{
						const float* wl = t.wL + x * kLanczosStride; const float* wr = t.wR + x * kLanczosStride;
						iL = lanczosAt(csf, t.baseL[x], wl, N); iR = lanczosAt(csf, t.baseR[x], wr, N);
						const int np = t.phaseNormPos[x];
						mL = make_float2(0.f, 0.f); mR = make_float2(0.f, 0.f);
#pragma unroll
						for (int k = 0; k < kLanczosTaps; ++k)
						{
							const float2 a = phaseView(csf, t.baseL[x] + k, N, np), b = phaseView(csf, t.baseR[x] + k, N, np);
							mL.x += a.x * wl[k]; mL.y += a.y * wl[k];
							mR.x += b.x * wr[k]; mR.y += b.y * wr[k];
						}
					}
					const float cancellation = invSize * cabs(cadd(iL, iR));
					const float midRaw = invSize * (cabs(iL) + cabs(iR));
					cancelOut = 1.f - (midRaw > 0.f ? cancellation / midRaw : 0.f);
					mid = invSize * (cabs(mL) + cabs(mR));
				}
				else
				{
					const int np = t.phaseNormFinal;
					int maxBin = 0; float maxValue = 0.f;
					for (int c = 0; c < bm.z; ++c)
					{
						const int off = bm.y + c;
						const float m = fmaxf(cnorm(phaseView(csf, off, N, np)), cnorm(phaseView(csf, N - off, N, np)));
						if (m > maxValue) { maxValue = m; maxBin = off; }
					}
					const float2 l = phaseView(csf, maxBin, N, np), r = phaseView(csf, N - maxBin, N, np);
					const float interference = invSize * cabs(cadd(l, r));
					mid = invSize * (cabs(l) + cabs(r));
					cancelOut = 1.f - (mid > 0.f ? interference / mid : 0.f);
				}
				work[x] = make_float2(mid, cancelOut); // wsp[2x], wsp[2x+1]
				continue;
			}

			float2 vL, vR = make_float2(0.f, 0.f);
			if (bm.x == 0)
			{
				if (t.binpol == SGZ_BIN_NONE) { vL = csf[bm.y]; if (chans == 2) vR = csf[N - bm.y]; }
				else if (t.binpol == SGZ_BIN_LINEAR) { vL = linearAt(csf, t.posL[x], N); if (chans == 2) vR = linearAt(csf, t.posR[x], N); }
				else
				{
					vL = lanczosAt(csf, t.baseL[x], t.wL + x * kLanczosStride, N);
					if (chans == 2) vR = lanczosAt(csf, t.baseR[x], t.wR + x * kLanczosStride, N);
				}
			}
			else
			{
				int maxL = bm.w, maxR = bm.w; float bestL = 0.f, bestR = 0.f;
				for (int c = 0; c < bm.z; ++c)
				{
					const int off = bm.y + c;
					const float mL = cnorm(csf[off]);
					if (mL > bestL) { bestL = mL; maxL = off; }
					if (chans == 2)
					{
						const float mR = cnorm(csf[N - off]);
						if (mR > bestR) { bestR = mR; maxR = N - off; }
					}
				}
				vL = csf[maxL];
				if (chans == 2) vR = csf[maxR];
			}
			work[x] = make_float2(invSize * vL.x, invSize * vL.y);
			if (chans == 2) work[P + x] = make_float2(invSize * vR.x, invSize * vR.y);
		}
	}

	// ------------------------------------------------------------------------------------------
	// staged kernel 4: mapAndTransformDFTFilters (TransformDSP.inl:1299-1435)
	// states / results: [stream][graph][P] float2
	// ------------------------------------------------------------------------------------------
	__device__ __forceinline__ float logScale(float slope, float st, float minFracRecip, float dy, float clip)
	{
		const float d = slope * st * minFracRecip;
		return d > 0.f ? logf(d) * dy : clip;
	}

	__global__ void k_post(DeviceTables t, const float2* __restrict__ workAll, float2* __restrict__ states, float2* __restrict__ results)
	{
		const int x = blockIdx.x * blockDim.x + threadIdx.x;
		const int s = blockIdx.y;
		const int P = int(t.P);
		if (x >= P) return;
		const int conf = t.conf;
		const bool mono = isMonoConf(conf);
		const float2* work = workAll + uint64_t(s) * t.workStride;
		const float slope = t.slope[x];
		const float poles[2] = { t.pole0, t.pole1 };
		const float phasePoles[2] = { t.phasePole0, t.phasePole1 };
		if (mono || conf == SGZ_CH_COMPLEX)
		{
			const float2 v = work[x];
			const float mag = sqrtf(v.x * v.x + v.y * v.y);
#pragma unroll
			for (int k = 0; k < 2; ++k)
			{
				float2* st = states + (uint64_t(s) * 2 + k) * P + x;
				float2 sv = *st;
				sv.x *= poles[k];
				if (mag > sv.x) sv.x = mag;
				*st = sv;
				results[(uint64_t(s) * 2 + k) * P + x] = make_float2(logScale(slope, sv.x, t.minFracRecip, t.deltaYRecip, t.lowerClip), 0.f);
			}
		}
		else if (conf == SGZ_CH_PHASE)
		{
			const float2 v = work[x];
			const float mag = v.x * 0.5f;
			float ph = v.y;
#pragma unroll
			for (int k = 0; k < 2; ++k)
			{
				float2* st = states + (uint64_t(s) * 2 + k) * P + x;
				float2 sv = *st;
				sv.x *= poles[k];
				if (mag > sv.x) sv.x = mag;
				ph *= mag; // inside the k loop, as the reference (:1417)
				sv.y = ph + phasePoles[k] * (sv.y - ph);
				*st = sv;
				results[(uint64_t(s) * 2 + k) * P + x] = make_float2(
					logScale(slope, sv.x, t.minFracRecip, t.deltaYRecip, t.lowerClip),
					logScale(slope, sv.y, t.minFracRecip, t.deltaYRecip, t.lowerClip));
			}
		}
		else
		{
			const float2 l = work[x], r = work[P + x];
			const float lmag = sqrtf(l.x * l.x + l.y * l.y), rmag = sqrtf(r.x * r.x + r.y * r.y);
#pragma unroll
			for (int k = 0; k < 2; ++k)
			{
				float2* st = states + (uint64_t(s) * 2 + k) * P + x;
				float2 sv = *st;
				sv.x *= poles[k]; sv.y *= poles[k];
				if (lmag > sv.x) sv.x = lmag;
				if (rmag > sv.y) sv.y = rmag;
				*st = sv;
				results[(uint64_t(s) * 2 + k) * P + x] = make_float2(
					logScale(slope, sv.x, t.minFracRecip, t.deltaYRecip, t.lowerClip),
					logScale(slope, sv.y, t.minFracRecip, t.deltaYRecip, t.lowerClip));
			}
		}
	}

	// ------------------------------------------------------------------------------------------
	// spectrogram colour columns: blendAndDispatchSpectrums (SpectrumDSP.cpp:110-206).  uint8 output,
	// so the float arithmetic that decides it uses non-contractable round-to-nearest intrinsics.
	// ------------------------------------------------------------------------------------------
	struct ColourArgs
	{
		const float2* results; // [S][F][2][P]
		uint32_t nFrames, pairs, P;
		float ratios[6];
		const float* colours;  // [pairs][6][3]
		uchar4* rgba;          // [S / pairs][F][P]
	};

	__global__ void k_colour_columns(ColourArgs a)
	{
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t f = blockIdx.y, view = blockIdx.z;
		if (i >= a.P) return;
		float cb[3] = { 0.f, 0.f, 0.f };
		for (uint32_t p = 0; p < a.pairs; ++p)
		{
			const uint32_t s = view * a.pairs + p;
			const float intensity = a.results[((uint64_t(s) * a.nFrames + f) * 2 + 0) * a.P + i].x; // LineMain .magnitude
			if (intensity < 0) continue;
			const float* sca = a.colours + p * 18;
			float colour[3] = { 0.f, 0.f, 0.f };
			if (intensity < 0.999f)
			{
				float accumulatedSum = 0;
				for (int c = 1; c < 6; ++c)
				{
					const float nextScale = a.ratios[c];
					accumulatedSum = __fadd_rn(accumulatedSum, nextScale);
					if (accumulatedSum >= intensity)
					{
						const float mn = __fsub_rn(accumulatedSum, nextScale), mx = accumulatedSum;
						const float mix = __fdiv_rn(__fsub_rn(intensity, mn), __fsub_rn(mx, mn));
						const float imix = __fsub_rn(1.f, mix);
						for (int k = 0; k < 3; ++k)
							colour[k] = __fadd_rn(__fmul_rn(sca[(c - 1) * 3 + k], imix), __fmul_rn(sca[c * 3 + k], mix));
						break;
					}
				}
			}
			else
			{
				for (int k = 0; k < 3; ++k) colour[k] = sca[15 + k];
			}
			for (int k = 0; k < 3; ++k) cb[k] = __fadd_rn(cb[k], __fmul_rn(__fsub_rn(1.f, cb[k]), colour[k])); // GL_ONE_MINUS_SRC_COLOR
		}
		uchar4 px;
		px.x = static_cast<unsigned char>(__fmul_rn(cb[0], 255.f));
		px.y = static_cast<unsigned char>(__fmul_rn(cb[1], 255.f));
		px.z = static_cast<unsigned char>(__fmul_rn(cb[2], 255.f));
		px.w = 255;
		a.rgba[(uint64_t(view) * a.nFrames + f) * a.P + i] = px;
	}

#include "sgz_fused.cuh"
#include "sgz_fused8.cuh"
#include "sgz_resonator.cuh"
#include "sgz_resonator_tc.cuh"

} // namespace sgz

// ----------------------------------------------------------------------------------------------
// handle
// ----------------------------------------------------------------------------------------------
using namespace sgz;

struct sgz_spectrum
{
	sgz_spectrum_config cfg{};
	Tables tab;
	DeviceTables dev{};
	uint32_t nStreams{ 0 };
	int device{ 0 };
	cudaStream_t stream{ nullptr };
	bool ownStream{ false };
	std::vector<void*> allocations;
	// per-stream buffers
	float2* dCsf{ nullptr };     // [S][N+1]  audioMemory
	float2* dWork{ nullptr };    // [S][max(N, 2P)] workingMemory (FFT scratch / transform result)
	float2* dStates{ nullptr };  // [S][2][P]
	float2* dResults{ nullptr }; // [S][2][P]
	float2* dResState{ nullptr };   // [S][ch][V][P] resonator state (cresonator)
	float2* dResPartial{ nullptr }; size_t resPartialBytes{ 0 }; // time-slice scratch of k_resonate
	float* dResTcA{ nullptr };      // tensor-core resonator: pole-power operand tiles (sgz_resonator_tc.cuh)
	float2* dResTcPowL{ nullptr };  // ... and p^L per filter row
	int resonatorPath{ SGZ_RESONATOR_AUTO };
	// staging for host-resident callers
	float* dAudioStage{ nullptr }; size_t audioStageBytes{ 0 };
	float* dResultStage{ nullptr }; size_t resultStageBytes{ 0 };
	int stage{ 0 }; // 0 idle, 1 prepared, 2 transformed, 3 mapped
	uint64_t processedSamplesSinceLastFrame{ 0 }; // TransformPair.h:189 (one authority counter for all pairs, SpectrumDSP.cpp:81,93)
	int fusedK1{ 0 }; // first-pass radix of the fused kernel k_fused8<K1> this handle uses (0: staged kernels only)
	cudaStream_t copyIn{ nullptr }, copyOut{ nullptr }; // host-buffer pipeline of process_frames
	std::vector<cudaEvent_t> events;
	uint32_t pipeChunks{ 16 };    // stream chunks of the host-buffer pipeline (measured on BASELINE config 3: 8 -> 2.59 M, 16 -> 2.68 M, 32 -> 2.66 M, 64 -> 2.53 M frames/s)
	bool seamsOk{ false };        // the unchained launch mode is exact for this configuration (monotone hold and display scale)
	uint32_t fusedCapacity{ 0 };  // CTAs of this handle's fused kernel the device holds at once (the launch size of k_fused8)
	uint32_t fusedEpoch{ 0 };     // launch counter: the value a seam flag must show (no clearing between launches)
	float2* dSeam{ nullptr };     // [2 capacity][2][P] peak-hold states at the seams between CTA ranges
	uint32_t* dSeamFlags{ nullptr }; // [capacity]
	float2* dSpare{ nullptr };    // [P] sink for results an output mode does not keep (FusedArgs::spare)
	void* dTrack{ nullptr }; // frequency tracker: [S] bounds in, [S] records out
	// device-resident audio history (sgz_spectrum_history_*): per stream a ring of `ringCapacity` stereo
	// samples stored twice back to back ([S][2 C] float2), so that any run of <= C samples ending at the
	// write head is one contiguous span whatever the wrap position
	float2* dRing{ nullptr };
	uint64_t ringCapacity{ 0 };
	uint64_t ringTotal{ 0 };      // samples appended since reserve (absolute time of the write head)
	uint64_t ringLastFrame{ 0 };  // absolute time of the last frame boundary consumed by history_process
};

namespace
{
	struct DeviceGuard
	{
		int prev{ -1 };
		bool ok{ true };
		explicit DeviceGuard(int dev)
		{
			if (dev < 0) { ok = false; return; }
			if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
			if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
		}
		~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
	};

	template<typename T>
	int upload(sgz_spectrum* h, const std::vector<T>& src, const T** dst)
	{
		void* p = nullptr;
		const size_t bytes = std::max<size_t>(src.size(), 1) * sizeof(T);
		SGZ_CUDA(cudaMalloc(&p, bytes));
		h->allocations.push_back(p);
		if (!src.empty()) SGZ_CUDA(cudaMemcpy(p, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
		*dst = static_cast<const T*>(p);
		return SGZ_OK;
	}

	template<typename T>
	int allocZero(sgz_spectrum* h, size_t count, T** dst)
	{
		void* p = nullptr;
		SGZ_CUDA(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
		h->allocations.push_back(p);
		SGZ_CUDA(cudaMemset(p, 0, std::max<size_t>(count, 1) * sizeof(T)));
		*dst = static_cast<T*>(p);
		return SGZ_OK;
	}

	std::vector<float2> twiddleTable(size_t count, double denom, const std::vector<uint64_t>& exponents)
	{
		std::vector<float2> out(count);
		const double twoPi = 6.283185307179586476925286766559;
		for (size_t i = 0; i < count; ++i)
		{
			const double a = -twoPi * double(exponents[i] % uint64_t(denom)) / denom;
			out[i] = make_float2(float(std::cos(a)), float(std::sin(a)));
		}
		return out;
	}

	int ensureStage(sgz_spectrum* h, float** buf, size_t* have, size_t need)
	{
		if (*have >= need) return SGZ_OK;
		if (*buf) { cudaFree(*buf); *buf = nullptr; *have = 0; }
		SGZ_CUDA(cudaMalloc(reinterpret_cast<void**>(buf), need));
		*have = need;
		return SGZ_OK;
	}

	// FusedLaunch: what one k_fused8 launch of `a` should look like.  query = true only fills h->fusedCapacity (the
	// number of CTAs of this handle's kernel the device holds at once), launching nothing.
	struct FusedPlan { uint32_t ctas; bool chained; bool cooperative; bool dual = false; }; // dual: pairs of ranges run as the two halves of one 1024-thread CTA

	template<void (*kern)(DeviceTables, FusedArgs)>
	int launchFusedKernel(sgz_spectrum* h, int threads, size_t smem, FusedArgs& a, const FusedPlan* plan)
	{
		// the opt-in shared-memory limit belongs to the kernel function on a device, not to a handle: keep a running
		// maximum per (kernel, device) and only ever raise it, so handles with different display-point counts coexist
		// (the limit is kept no larger than needed: the driver sizes the shared-memory carve-out, and with it what is
		// left for L1, from it)
		static std::atomic<size_t> attrBytes[64];
		std::atomic<size_t>& limit = attrBytes[h->device & 63];
		if (limit.load(std::memory_order_acquire) < smem)
		{
			static std::mutex attrMutex;
			std::lock_guard<std::mutex> lock(attrMutex);
			if (limit.load(std::memory_order_relaxed) < smem)
			{
				SGZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
#ifdef SGZ8_CARVEOUT
				SGZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, SGZ8_CARVEOUT));
#endif
				limit.store(smem, std::memory_order_release);
			}
		}
		if (!plan)
		{
			int perSm = 0, sms = 0;
			SGZ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kern, threads, smem));
			SGZ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
			h->fusedCapacity = uint32_t(std::max(1, perSm) * std::max(1, sms));
#ifdef SGZ_DIAG_CAPACITY_DIV // timing experiments: fewer CTAs than the GPU holds (2: one per SM for the 8192-point kernel)
			h->fusedCapacity = std::max<uint32_t>(1, h->fusedCapacity / SGZ_DIAG_CAPACITY_DIV);
#endif
			return SGZ_OK;
		}
		a.chained = plan->chained ? 1u : 0u;
		a.epoch = ++h->fusedEpoch;
		a.seam = h->dSeam; a.flags = h->dSeamFlags;
		if (plan->cooperative)
		{
			// CTAs wait on one another's seam flags: the launch must be co-resident
			void* params[] = { &h->dev, &a };
			SGZ_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(kern), dim3(plan->ctas), dim3(threads), params, smem, h->stream));
		}
		else kern<<<plan->ctas, threads, smem, h->stream>>>(h->dev, a);
		SGZ_CUDA(cudaGetLastError());
		return SGZ_OK;
	}

	template<int K1, bool STEREO, int MATRIX>
	int launchFused8Binpol(sgz_spectrum* h, FusedArgs& a, const FusedPlan* plan)
	{
		const size_t smem = Fused8GeomT<K1>::smemBytes(h->tab.P);
		constexpr int T = Fused8GeomT<K1>::T;
		switch (h->cfg.bin_polation)
		{
#ifndef SGZ_HEADLINE_ONLY
		case SGZ_BIN_NONE: return launchFusedKernel<k_fused8<K1, STEREO, SGZ_BIN_NONE, MATRIX>>(h, T, smem, a, plan);
		case SGZ_BIN_LINEAR: return launchFusedKernel<k_fused8<K1, STEREO, SGZ_BIN_LINEAR, MATRIX>>(h, T, smem, a, plan);
#endif
		default:
			if constexpr (K1 == 8 && STEREO && MATRIX == 0)
				if (plan && plan->dual)
				{
					FusedPlan halves = *plan;
					halves.ctas = plan->ctas / 2;
					return launchFusedKernel<k_fused8<K1, STEREO, SGZ_BIN_LANCZOS, MATRIX, true>>(h, 2 * T, 2 * ((smem + 127) & ~size_t(127)), a, &halves);
				}
			return launchFusedKernel<k_fused8<K1, STEREO, SGZ_BIN_LANCZOS, MATRIX>>(h, T, smem, a, plan);
		}
	}

	template<int K1>
	int launchFusedK1(sgz_spectrum* h, FusedArgs& a, const FusedPlan* plan)
	{
		if (h->cfg.configuration == SGZ_CH_SEPARATE) return launchFused8Binpol<K1, true, 0>(h, a, plan);
#ifdef SGZ_HEADLINE_ONLY // (A/B builds of the bench configuration: nothing else is instantiated)
		return fail(SGZ_EUNSUPPORTED, "headline-only build");
#endif
		if (h->cfg.configuration == SGZ_CH_COMPLEX) return launchFused8Binpol<K1, false, 2>(h, a, plan);
		if (h->cfg.configuration == SGZ_CH_PHASE) return launchFused8Binpol<K1, true, 3>(h, a, plan);
		return h->tab.stateChannels == 2 ? launchFused8Binpol<K1, true, 1>(h, a, plan) : launchFused8Binpol<K1, false, 1>(h, a, plan);
	}

	// plan = nullptr: capacity query only (sgz_spectrum_create)
	int launchFused(sgz_spectrum* h, FusedArgs& a, const FusedPlan* plan)
	{
		switch (h->fusedK1)
		{
#ifndef SGZ_HEADLINE_ONLY
		case 16: return launchFusedK1<16>(h, a, plan);
		case 4: return launchFusedK1<4>(h, a, plan);
		case 2: return launchFusedK1<2>(h, a, plan);
		case 1: return launchFusedK1<1>(h, a, plan);
#endif
		case 8: return launchFusedK1<8>(h, a, plan);
		default: return fail(SGZ_EUNSUPPORTED, "no fused kernel for this configuration");
		}
	}

	size_t audioFloatsPerStream(uint64_t nSamples) { return size_t(nSamples) * 2; }

	// copies caller audio to the device if needed; returns the device pointer to use
	int stageAudio(sgz_spectrum* h, const float* audio, int mem, uint64_t nSamples, uint64_t streamStride, const float** dptr, float** slot, size_t* slotBytes)
	{
		if (mem == SGZ_MEM_DEVICE) { *dptr = audio; return SGZ_OK; }
		const size_t floats = size_t(streamStride) * (h->nStreams - 1) + audioFloatsPerStream(nSamples);
		const int rc = ensureStage(h, slot, slotBytes, floats * sizeof(float));
		if (rc != SGZ_OK) return rc;
		SGZ_CUDA(cudaMemcpyAsync(*slot, audio, floats * sizeof(float), cudaMemcpyHostToDevice, h->stream));
		*dptr = *slot;
		return SGZ_OK;
	}

}

#define SGZ_GUARD(h)                                                                                        \
	DeviceGuard g__((h)->device);                                                                           \
	if (!g__.ok) return fail(SGZ_ECUDA, (h)->device < 0 ? "tables-only handle (SGZ_DEVICE_NONE): no compute, and there is no CPU fallback" : "cudaSetDevice failed")

// ----------------------------------------------------------------------------------------------
// C ABI
// ----------------------------------------------------------------------------------------------
extern "C"
{
	const char* sgz_last_error(void) { return g_lastError.c_str(); }

	void sgz_spectrum_default_config(sgz_spectrum_config* cfg)
	{
		if (!cfg) return;
		std::memset(cfg, 0, sizeof(*cfg));
		cfg->abi_version = SGZ_ABI_VERSION;
		cfg->sample_rate = 48000.f;
		cfg->window_size = 8192;
		cfg->axis_points = 1024;
		cfg->algo = SGZ_ALGO_FFT;
		cfg->configuration = SGZ_CH_SEPARATE;
		cfg->bin_polation = SGZ_BIN_LANCZOS;
		cfg->display_mode = SGZ_DISPLAY_LINEGRAPH;
		cfg->view_scaling = SGZ_VIEW_LOGARITHMIC;
		cfg->view_left = 0.0; cfg->view_right = 1.0;
		cfg->min_freq = 10.0;
		cfg->slope_a = 0.0; cfg->slope_b = 1.0;
		cfg->low_dbs = -120.0; cfg->high_dbs = 0.0; cfg->clip_db = -384.0;
		cfg->pole[0] = 0.95f; cfg->pole[1] = 0.99f;
		cfg->window_type = SGZ_WIN_BLACKMAN_HARRIS;
		cfg->window_shape = SGZ_WINSHAPE_PERIODIC;
		cfg->sample_buffer_size = 200;
		cfg->free_q = 0;
	}

	void sgz_spectrum_destroy(sgz_spectrum* h)
	{
		if (!h) return;
		if (h->device != SGZ_DEVICE_NONE)
		{
			DeviceGuard g(h->device);
			if (h->stream) cudaStreamSynchronize(h->stream);
			for (void* p : h->allocations) cudaFree(p);
			if (h->dRing) cudaFree(h->dRing);
			if (h->dTrack) cudaFree(h->dTrack);
			if (h->dAudioStage) cudaFree(h->dAudioStage);
			if (h->dResultStage) cudaFree(h->dResultStage);
			if (h->dResPartial) cudaFree(h->dResPartial);
			for (cudaEvent_t e : h->events) cudaEventDestroy(e);
			if (h->copyIn) cudaStreamDestroy(h->copyIn);
			if (h->copyOut) cudaStreamDestroy(h->copyOut);
			if (h->ownStream && h->stream) cudaStreamDestroy(h->stream);
		}
		delete h;
	}

	int sgz_spectrum_create(const sgz_spectrum_config* cfg, uint32_t n_streams, int device, sgz_spectrum** out)
	{
		if (!cfg || !out || n_streams == 0) return fail(SGZ_EINVAL, "null argument or zero streams");
		*out = nullptr;
		if (device == SGZ_DEVICE_NONE)
		{
			// tables-only handle: the derived constants can be inspected, nothing can be computed
			sgz_spectrum* h = new (std::nothrow) sgz_spectrum();
			if (!h) return fail(SGZ_ENOMEM, "out of host memory");
			h->cfg = *cfg; h->nStreams = n_streams; h->device = SGZ_DEVICE_NONE;
			const int rc = buildTables(*cfg, h->tab);
			if (rc != SGZ_OK) { delete h; return fail(rc, "invalid spectrum configuration"); }
			h->fusedK1 = (h->tab.fusedOk && fusedSharedMemoryFits(h->tab.N, h->tab.P)) ? fusedK1For(h->tab.N) : 0;
			*out = h;
			return SGZ_OK;
		}
		int count = 0;
		if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return fail(SGZ_ECUDA, "no CUDA device (this library has no CPU fallback)");
		if (device < 0 || device >= count) return fail(SGZ_EINVAL, "device %d out of range (%d devices)", device, count);

		sgz_spectrum* h = new (std::nothrow) sgz_spectrum();
		if (!h) return fail(SGZ_ENOMEM, "out of host memory");
		h->cfg = *cfg; h->nStreams = n_streams; h->device = device;
		int rc = buildTables(*cfg, h->tab);
		if (rc != SGZ_OK) { delete h; return fail(rc, "invalid spectrum configuration"); }

		DeviceGuard guard(device);
		if (!guard.ok) { delete h; return fail(SGZ_ECUDA, "cudaSetDevice(%d) failed", device); }
		auto bail = [&](int code) { sgz_spectrum_destroy(h); return code; };
		if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(fail(SGZ_ECUDA, "cudaStreamCreate failed"));
		h->ownStream = true;

		const Tables& t = h->tab;
		DeviceTables& d = h->dev;
		d.N = t.N; d.W = t.W; d.P = t.P; d.M = t.numBins; d.linLimit = t.linLimit;
		d.conf = cfg->configuration; d.binpol = cfg->bin_polation;
		d.deltaYLn2 = float(double(t.deltaYRecip) * 0.69314718055994530942);
		d.invSize = t.invSize; d.deltaYRecip = t.deltaYRecip; d.minFracRecip = t.minFracRecip; d.lowerClip = t.lowerClip;
		d.pole0 = t.pole[0]; d.pole1 = t.pole[1]; d.phasePole0 = t.phasePole[0]; d.phasePole1 = t.phasePole[1];
		d.phaseNormFinal = t.phaseNormFinal;
		d.safeFrom = t.safeFrom;
		d.workStride = uint32_t(std::max<size_t>(t.N, size_t(2) * t.P));
		d.winPolyOk = t.winPolyOk ? 1 : 0;
		for (int k = 0; k < 5; ++k) d.winPoly[k] = t.winPoly[k];
		d.winStep = make_float2(float(std::cos(6.283185307179586476925286766559 / double(t.N))), float(std::sin(6.283185307179586476925286766559 / double(t.N))));

		std::vector<int4> binmap(t.P);
		for (uint32_t x = 0; x < t.P; ++x) binmap[x] = make_int4(t.mode[x], t.first[x], t.count[x], t.fallback[x]);
#define SGZ_TRY(expr) do { rc = (expr); if (rc != SGZ_OK) return bail(rc); } while (0)
		SGZ_TRY(upload(h, t.window, &d.window));
		SGZ_TRY(upload(h, t.slope, &d.slope));
		SGZ_TRY(upload(h, binmap, &d.binmap));
		SGZ_TRY(upload(h, t.posL, &d.posL));
		SGZ_TRY(upload(h, t.posR, &d.posR));
		SGZ_TRY(upload(h, t.baseL, &d.baseL));
		SGZ_TRY(upload(h, t.baseR, &d.baseR));
		SGZ_TRY(upload(h, t.wL, &d.wL));
		SGZ_TRY(upload(h, t.wR, &d.wR));
		{
			// fused kernel: word r of point x at [r][x], so a warp's 32 points are 512 contiguous bytes per load
			std::vector<float4> l4(size_t(3) * t.P, make_float4(0.f, 0.f, 0.f, 0.f)), r4(l4);
			if (t.wL.size() >= size_t(t.P) * kLanczosStride)
				for (uint32_t x = 0; x < t.P; ++x)
					for (int r = 0; r < 3; ++r)
					{
						const float* wl = &t.wL[size_t(x) * kLanczosStride + 4 * r];
						l4[size_t(r) * t.P + x] = make_float4(wl[0], wl[1], wl[2], wl[3]);
						if (t.wR.size() >= size_t(t.P) * kLanczosStride)
						{
							const float* wr = &t.wR[size_t(x) * kLanczosStride + 4 * r];
							r4[size_t(r) * t.P + x] = make_float4(wr[0], wr[1], wr[2], wr[3]);
						}
					}
			SGZ_TRY(upload(h, l4, &d.wL4));
			SGZ_TRY(upload(h, r4, &d.wR4));
		}
		SGZ_TRY(upload(h, t.phaseNormPos, &d.phaseNormPos));
		{
			std::vector<float2> coeff(t.resCoeffReal.size());
			for (size_t k = 0; k < coeff.size(); ++k) coeff[k] = make_float2(t.resCoeffReal[k], t.resCoeffImag[k]);
			SGZ_TRY(upload(h, coeff, &d.resCoeff));
			SGZ_TRY(upload(h, t.resGain, &d.resGain));
			SGZ_TRY(upload(h, t.resWindowTaps, &d.resTaps));
			d.resVectors = t.resonatorVectors; d.resChannels = t.stateChannels;
		}
		{
			std::vector<uint64_t> e(t.N / 2);
			for (size_t k = 0; k < e.size(); ++k) e[k] = k;
			SGZ_TRY(upload(h, twiddleTable(e.size(), double(t.N), e), &d.twN));
		}
		if (t.N == 1024 || t.N == 2048 || t.N == 4096 || t.N >= 8192)
		{
			// tables of the radix-8 register FFT (k_fused8 and the staged k_fft8): W_n^k for the last radix-2, and
			// [8][8] W_64^(ab) | [8][64] W_M^(v k1) | [8][64] W_512^(v k2); n = N, or 16384 for the sub-transforms of a larger N
			const uint32_t subN = std::min<uint32_t>(t.N, 16384), M = subN / 2;
			std::vector<uint64_t> e3(M);
			for (uint32_t k = 0; k < M; ++k) e3[k] = k;
			SGZ_TRY(upload(h, twiddleTable(e3.size(), double(subN), e3), &d.tw3));
			std::vector<uint64_t> eA(64), eB(512), eC(512);
			for (int x = 0; x < 8; ++x) for (int y = 0; y < 8; ++y) eA[x * 8 + y] = uint64_t(x) * y;
			for (int k = 0; k < 8; ++k) for (int v = 0; v < 64; ++v) { eB[k * 64 + v] = uint64_t(v) * k; eC[k * 64 + v] = uint64_t(v) * k; }
			std::vector<float2> tw8;
			const auto A = twiddleTable(64, 64.0, eA), B = twiddleTable(512, 4096.0, eB), C = twiddleTable(512, 512.0, eC);
			tw8.insert(tw8.end(), A.begin(), A.end()); tw8.insert(tw8.end(), B.begin(), B.end()); tw8.insert(tw8.end(), C.begin(), C.end());
			SGZ_TRY(upload(h, tw8, &d.tw8));
		}
		h->fusedK1 = (t.fusedOk && fusedSharedMemoryFits(t.N, t.P)) ? fusedK1For(t.N) : 0;
		const size_t workPerStream = std::max<size_t>(t.N, size_t(2) * t.P);
		SGZ_TRY(allocZero(h, size_t(n_streams) * (t.N + 1), &h->dCsf));
		SGZ_TRY(allocZero(h, size_t(n_streams) * workPerStream, &h->dWork));
		SGZ_TRY(allocZero(h, size_t(n_streams) * 2 * t.P, &h->dStates));
		SGZ_TRY(allocZero(h, size_t(n_streams) * 2 * t.P, &h->dResults));
		if (cfg->algo == SGZ_ALGO_RSNT) SGZ_TRY(allocZero(h, size_t(n_streams) * t.stateChannels * t.resonatorVectors * t.P, &h->dResState));
		if (cfg->algo == SGZ_ALGO_RSNT)
		{
			// operands of k_resonate_tc: for every filter row g = z P + i the powers p^(L-1-k), k = 0 .. L-1, in double,
			// split into a TF32 head and tail, stored per 128-row tile in the kernel's shared-memory order
			const uint32_t half = t.P * t.resonatorVectors, rowTiles = (half + kTcRows - 1) / kTcRows;
			std::vector<float> aTiles(size_t(rowTiles) * 4 * kTcAFloats, 0.f);
			std::vector<float2> powL(size_t(rowTiles) * kTcRows, make_float2(0.f, 0.f));
			auto tf32 = [](double v) { // round to 11 significant bits (ties away), like cvt.rna.tf32.f32
				float f = float(v); uint32_t u; std::memcpy(&u, &f, 4);
				u = (u + 0x1000u) & 0xFFFFE000u; std::memcpy(&f, &u, 4); return f; };
			for (uint32_t g = 0; g < half; ++g)
			{
				const std::complex<double> p(t.resCoeffReal[g], t.resCoeffImag[g]);
				std::complex<double> pw(1.0, 0.0);
				float* tile = aTiles.data() + size_t(g / kTcRows) * 4 * kTcAFloats;
				const uint32_t row = g % kTcRows, rg = row / 8, r = row % 8;
				for (int e = 0; e < kTcL; ++e, pw *= p)
				{
					const int k = kTcL - 1 - e;   // x[k] is multiplied by p^(L-1-k) = p^e
					const size_t off = ((size_t(k / 4) * (kTcRows / 8) + rg) * 8 + r) * 4 + (k % 4);
					const float reH = tf32(pw.real()), imH = tf32(pw.imag());
					tile[0 * kTcAFloats + off] = reH; tile[1 * kTcAFloats + off] = tf32(pw.real() - reH);
					tile[2 * kTcAFloats + off] = imH; tile[3 * kTcAFloats + off] = tf32(pw.imag() - imH);
				}
				powL[g] = make_float2(float(pw.real()), float(pw.imag())); // p^L
			}
			const float* da = nullptr; const float2* dp = nullptr;
			SGZ_TRY(upload(h, aTiles, &da));
			SGZ_TRY(upload(h, powL, &dp));
			h->dResTcA = const_cast<float*>(da); h->dResTcPowL = const_cast<float2*>(dp);
			if (cudaFuncSetAttribute(k_resonate_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kTcSmemBytes)) != cudaSuccess)
				return bail(fail(SGZ_ECUDA, "cudaFuncSetAttribute(k_resonate_tc) failed"));
		}
		if (h->fusedK1 != 0)
		{
			FusedArgs none{};
			SGZ_TRY(launchFused(h, none, nullptr)); // capacity query only
			SGZ_TRY(allocZero(h, size_t(2) * h->fusedCapacity * 2 * t.P, &h->dSeam));
			SGZ_TRY(allocZero(h, size_t(h->fusedCapacity), &h->dSeamFlags));
			SGZ_TRY(allocZero(h, size_t(t.P), &h->dSpare));
			// carrying the hold across zero-started parts relies on x -> x * pole and on the dB scale being monotone
			// non-decreasing: non-negative poles and slope, an upright dB range, and a clip value below every reachable one
			h->seamsOk = cfg->pole[0] >= 0.f && cfg->pole[1] >= 0.f && cfg->slope_b >= 0.0 && d.deltaYLn2 > 0.f
				&& d.lowerClip <= -104.f * d.deltaYRecip;
		}
		// the tables and cleared buffers above went through the legacy stream; the handle's own stream is non-blocking
		if (cudaDeviceSynchronize() != cudaSuccess) return bail(fail(SGZ_ECUDA, "cudaDeviceSynchronize failed in create"));
#undef SGZ_TRY
		*out = h;
		return SGZ_OK;
	}

	int sgz_spectrum_get_info(const sgz_spectrum* h, sgz_spectrum_info* out)
	{
		if (!h || !out) return fail(SGZ_EINVAL, "null argument");
		std::memset(out, 0, sizeof(*out));
		out->transform_size = h->tab.N; out->window_size = h->tab.W; out->axis_points = h->tab.P;
		out->n_streams = h->nStreams; out->state_channels = h->tab.stateChannels;
		out->break_point = h->tab.P;
		for (uint32_t x = 0; x < h->tab.P; ++x) if (h->tab.mode[x] == 1) { out->break_point = x; break; }
		out->window_kernel_scale = h->tab.windowKernelScale; out->inv_size = h->tab.invSize; out->freq_to_bin = h->tab.freqToBin;
		out->resonator_vectors = h->tab.resonatorVectors; out->device = h->device; out->fused_available = h->fusedK1 != 0;
		return SGZ_OK;
	}

	const char* sgz_spectrum_kernel_name(const sgz_spectrum* h)
	{
		// the instantiation launchFused picks for this handle, spelled like the profiler spells it
		static thread_local char name[96];
		if (!h || h->fusedK1 == 0) return "";
		const int conf = h->cfg.configuration;
		const bool stereo = conf == SGZ_CH_SEPARATE || conf == SGZ_CH_PHASE || (conf != SGZ_CH_COMPLEX && h->tab.stateChannels == 2);
		const int matrix = conf == SGZ_CH_SEPARATE ? 0 : conf == SGZ_CH_COMPLEX ? 2 : conf == SGZ_CH_PHASE ? 3 : 1;
		std::snprintf(name, sizeof(name), "sgz::k_fused8<%d, %d, %d, %d, 0>", h->fusedK1, stereo ? 1 : 0, int(h->cfg.bin_polation), matrix);
		return name;
	}

	int sgz_spectrum_set_cuda_stream(sgz_spectrum* h, void* cuda_stream)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		SGZ_GUARD(h);
		if (h->stream) cudaStreamSynchronize(h->stream);
		if (h->ownStream && h->stream) cudaStreamDestroy(h->stream);
		h->stream = static_cast<cudaStream_t>(cuda_stream);
		h->ownStream = false;
		return SGZ_OK;
	}

	int sgz_spectrum_synchronize(sgz_spectrum* h)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		SGZ_GUARD(h);
		SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	int sgz_spectrum_get_window_kernel(const sgz_spectrum* h, float* out, size_t n)
	{
		if (!h || !out || n != h->tab.N) return fail(SGZ_EINVAL, "window kernel has %u entries", h ? h->tab.N : 0);
		std::memcpy(out, h->tab.window.data(), n * sizeof(float));
		return SGZ_OK;
	}
	int sgz_spectrum_get_mapped_frequencies(const sgz_spectrum* h, float* out, size_t n)
	{
		if (!h || !out || n != h->tab.P) return fail(SGZ_EINVAL, "mapped frequencies have %u entries", h ? h->tab.P : 0);
		std::memcpy(out, h->tab.mappedFreq.data(), n * sizeof(float));
		return SGZ_OK;
	}
	int sgz_spectrum_get_slope_map(const sgz_spectrum* h, float* out, size_t n)
	{
		if (!h || !out || n != h->tab.P) return fail(SGZ_EINVAL, "slope map has %u entries", h ? h->tab.P : 0);
		std::memcpy(out, h->tab.slope.data(), n * sizeof(float));
		return SGZ_OK;
	}
	int sgz_spectrum_get_bin_map(const sgz_spectrum* h, int32_t* mode, int32_t* first, int32_t* count, int32_t* fallback, size_t n)
	{
		if (!h || !mode || !first || !count || !fallback || n != h->tab.P) return fail(SGZ_EINVAL, "bin map has %u entries", h ? h->tab.P : 0);
		std::memcpy(mode, h->tab.mode.data(), n * sizeof(int32_t));
		std::memcpy(first, h->tab.first.data(), n * sizeof(int32_t));
		std::memcpy(count, h->tab.count.data(), n * sizeof(int32_t));
		std::memcpy(fallback, h->tab.fallback.data(), n * sizeof(int32_t));
		return SGZ_OK;
	}

	// nPrelim: samples of the preliminary buffer that count as the newest audio; prelimLength: samples per channel the
	// buffer actually holds (>= nPrelim) - the planar channel stride and what a host buffer needs staged
	static int prepareCommon(sgz_spectrum* h, const float* hist, uint64_t nHist, uint64_t histStride, const float* prelim, uint64_t nPrelim,
		uint64_t prelimStride, int mem, int layout, bool withPrelim, uint64_t prelimLength = 0)
	{
		if (prelimLength < nPrelim) prelimLength = nPrelim;
		if (!h || !hist) return fail(SGZ_EINVAL, "null argument");
		if (h->cfg.algo != SGZ_ALGO_FFT) return nHist < h->tab.W ? SGZ_SKIPPED : SGZ_OK; // the switch at TransformDSP.inl:51-53 only has an FFT case
		if (layout != SGZ_LAYOUT_PLANAR && layout != SGZ_LAYOUT_INTERLEAVED) return fail(SGZ_EINVAL, "bad layout");
		const uint32_t W = h->tab.W;
		if (nHist < W) { h->stage = 0; return SGZ_SKIPPED; } // TransformDSP.inl:45-46
		if (histStride < 2 * nHist && h->nStreams > 1) return fail(SGZ_EINVAL, "stream_stride smaller than one stream");
		if (layout == SGZ_LAYOUT_INTERLEAVED && ((histStride & 1) || (reinterpret_cast<uintptr_t>(hist) & 7) || (withPrelim && ((prelimStride & 1) || (reinterpret_cast<uintptr_t>(prelim) & 7)))))
			return fail(SGZ_EINVAL, "interleaved audio must be 8-byte aligned with an even stream_stride");
		SGZ_GUARD(h);
		PrepareArgs a{};
		a.layout = layout;
		int rc = stageAudio(h, hist, mem, nHist, histStride, &a.hist, &h->dAudioStage, &h->audioStageBytes);
		if (rc != SGZ_OK) return rc;
		a.nHist = nHist; a.histStride = histStride;
		if (!withPrelim)
		{
			a.histSkip = uint32_t(nHist - W); a.histCount = W; a.prelimCount = 0;
		}
		else
		{
			// TransformDSP.inl:245-257: stop = min(numSamples, W); offset = stop + extraDiscarded;
			// sizeToStopAt = W - offset.  Left/Right skip `offset` history samples; every other
			// configuration advances only an unused iterator and so starts at history sample 0.
			if (!prelim) return fail(SGZ_EINVAL, "null preliminary audio");
			const uint64_t extra = nHist - W;
			const uint64_t stop = std::min<uint64_t>(nPrelim, W);
			const uint64_t offset = stop + extra;
			if (offset > W) return fail(SGZ_EUNSUPPORTED, "history larger than the window together with preliminary audio overruns the reference's buffer (TransformDSP.inl:256-257)");
			const bool lr = h->cfg.configuration == SGZ_CH_LEFT || h->cfg.configuration == SGZ_CH_RIGHT;
			a.histSkip = lr ? uint32_t(offset) : 0u;
			a.histCount = uint32_t(W - offset);
			a.prelimCount = uint32_t(stop);
			if (mem == SGZ_MEM_DEVICE) a.prelim = prelim;
			else
			{
				const size_t floats = size_t(prelimStride) * (h->nStreams - 1) + audioFloatsPerStream(prelimLength);
				rc = ensureStage(h, &h->dResultStage, &h->resultStageBytes, floats * sizeof(float));
				if (rc != SGZ_OK) return rc;
				SGZ_CUDA(cudaMemcpyAsync(h->dResultStage, prelim, floats * sizeof(float), cudaMemcpyHostToDevice, h->stream));
				a.prelim = h->dResultStage;
			}
			a.nPrelim = prelimLength; a.prelimStride = prelimStride;
		}
		const dim3 block(256), grid((h->tab.N + 1 + 255) / 256, h->nStreams);
		k_prepare<<<grid, block, 0, h->stream>>>(h->dev, a, h->dCsf);
		SGZ_CUDA(cudaGetLastError());
		h->stage = 1;
		return SGZ_OK;
	}

	int sgz_spectrum_prepare(sgz_spectrum* h, const float* audio, int mem, int layout, uint64_t n_samples, uint64_t stream_stride)
	{
		return prepareCommon(h, audio, n_samples, stream_stride, nullptr, 0, 0, mem, layout, false);
	}

	int sgz_spectrum_prepare_preliminary(sgz_spectrum* h, const float* history, uint64_t n_history, uint64_t history_stride, const float* prelim,
		uint64_t n_prelim, uint64_t prelim_stride, int mem, int layout)
	{
		return prepareCommon(h, history, n_history, history_stride, prelim, n_prelim, prelim_stride, mem, layout, true);
	}

	int sgz_spectrum_transform(sgz_spectrum* h)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (h->cfg.algo != SGZ_ALGO_FFT) return SGZ_OK; // doTransform is a no-op for RSNT (:489-501)
		if (h->stage < 1) return fail(SGZ_ESTATE, "call prepareTransform() first (TransformPair.h:123-124)");
		SGZ_GUARD(h);
		const uint32_t N = h->tab.N;
#ifdef SGZ_STAGED_FFT_GENERIC // (development switch: every size through the shared-memory / global-pass kernels)
		const bool generic = true;
#else
		const bool generic = false;
#endif
		const uint64_t csfStride = uint64_t(N) + 1, workStride = std::max<size_t>(N, size_t(2) * h->tab.P);
#define SGZ_FFT8(K1, SRC, SSTRIDE, DST, DSTRIDE, SUBS)                                                                        \
		{                                                                                                                     \
			SGZ_CUDA(cudaFuncSetAttribute(k_fft8<K1>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Fft8Smem<K1>::bytes))); \
			k_fft8<K1><<<dim3(h->nStreams, SUBS), 64 * K1, Fft8Smem<K1>::bytes, h->stream>>>(h->dev, SRC, SSTRIDE, DST, DSTRIDE, SUBS); \
		}
		if (h->dev.tw8 && !generic && (N == 1024 || N == 2048 || N == 4096 || N == 8192 || N == 16384))
		{
			if (N == 16384) SGZ_FFT8(16, h->dCsf, csfStride, h->dCsf, csfStride, 1u)
			else if (N == 8192) SGZ_FFT8(8, h->dCsf, csfStride, h->dCsf, csfStride, 1u)
			else if (N == 4096) SGZ_FFT8(4, h->dCsf, csfStride, h->dCsf, csfStride, 1u)
			else if (N == 2048) SGZ_FFT8(2, h->dCsf, csfStride, h->dCsf, csfStride, 1u)
			else SGZ_FFT8(1, h->dCsf, csfStride, h->dCsf, csfStride, 1u)
		}
		else if (h->dev.tw8 && !generic && N > 16384 && N / 16384 <= 65535)
		{
			// N = subs * 16384: log2(subs) Stockham passes through global memory split the transform into `subs` interleaved
			// 16384-point ones, which the register FFT then finishes straight into audioMemory (out of place when an odd
			// number of passes left the data in the work buffer, in place otherwise): no copy back
			const uint32_t subs = N / 16384;
			int passes = 0; for (uint32_t v = subs; v > 1; v >>= 1) ++passes;
			const float2* src = h->dCsf; float2* dst = h->dWork;
			uint64_t sStride = csfStride, dStride = workStride;
			const dim3 block(256), grid((N / 2 + 255) / 256, h->nStreams);
			uint32_t n = N, st = 1;
			for (int pass = 0; pass < passes; ++pass, n >>= 1, st <<= 1)
			{
				k_fft_global_pass<<<grid, block, 0, h->stream>>>(h->dev.twN, src, dst, N, n, st, sStride, dStride);
				const float2* nsrc = dst; dst = const_cast<float2*>(src); src = nsrc;
				std::swap(sStride, dStride);
			}
			SGZ_FFT8(16, src, sStride, h->dCsf, csfStride, subs)
		}
#undef SGZ_FFT8
		else if (N <= 16384)
		{
			const size_t smem = size_t(N) * sizeof(float2);
			SGZ_CUDA(cudaFuncSetAttribute(k_fft_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
			k_fft_smem<<<h->nStreams, 256, smem, h->stream>>>(h->dev, h->dCsf);
		}
		else
		{
			const uint64_t csfStride = N + 1, workStride = std::max<size_t>(N, size_t(2) * h->tab.P);
			const float2* src = h->dCsf; float2* dst = h->dWork;
			uint64_t sStride = csfStride, dStride = workStride;
			const dim3 block(256), grid((N / 2 + 255) / 256, h->nStreams);
			for (uint32_t n = N, s = 1; n > 1; n >>= 1, s <<= 1)
			{
				k_fft_global_pass<<<grid, block, 0, h->stream>>>(h->dev.twN, src, dst, N, n, s, sStride, dStride);
				const float2* nsrc = dst; dst = const_cast<float2*>(src); src = nsrc;
				std::swap(sStride, dStride);
			}
			if (src != h->dCsf)
			{
				const dim3 g2((N + 255) / 256, h->nStreams);
				k_copy_strided<<<g2, block, 0, h->stream>>>(src, h->dCsf, N, sStride, csfStride);
			}
		}
		SGZ_CUDA(cudaGetLastError());
		h->stage = 2;
		return SGZ_OK;
	}

	int sgz_spectrum_map(sgz_spectrum* h)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (h->cfg.algo == SGZ_ALGO_RSNT)
		{
			// RSNT branch of mapToLinearSpace (TransformDSP.inl:1103-1133): windowed snapshot of the resonator state
			SGZ_GUARD(h);
			const dim3 block(256), grid((h->tab.P + 255) / 256, h->nStreams);
			k_resonator_snapshot<<<grid, block, 0, h->stream>>>(h->dev, h->dResState, h->dWork);
			SGZ_CUDA(cudaGetLastError());
			h->stage = 3;
			return SGZ_OK;
		}
		if (h->stage < 2) return fail(SGZ_ESTATE, "call doTransform() first (TransformPair.h:123-124)");
		if (h->tab.N < 3 || h->tab.sampleRate < 1) return SGZ_OK; // :519-520
		SGZ_GUARD(h);
		k_map<<<h->nStreams, 256, 0, h->stream>>>(h->dev, h->dCsf, h->dWork);
		SGZ_CUDA(cudaGetLastError());
		h->stage = 3;
		return SGZ_OK;
	}

	int sgz_spectrum_postprocess(sgz_spectrum* h)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (h->stage < 3) return fail(SGZ_ESTATE, "call mapToLinearSpace() first");
		SGZ_GUARD(h);
		const dim3 block(128), grid((h->tab.P + 127) / 128, h->nStreams);
		k_post<<<grid, block, 0, h->stream>>>(h->dev, h->dWork, h->dStates, h->dResults);
		SGZ_CUDA(cudaGetLastError());
		return SGZ_OK;
	}

	int sgz_spectrum_get_raw_fft(sgz_spectrum* h, float* out, int mem)
	{
		if (!h || !out) return fail(SGZ_EINVAL, "null argument");
		SGZ_GUARD(h);
		const uint32_t N = h->tab.N;
		SGZ_CUDA(cudaMemcpy2DAsync(out, size_t(N) * sizeof(float2), h->dCsf, size_t(N + 1) * sizeof(float2), size_t(N) * sizeof(float2), h->nStreams,
			mem == SGZ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
		if (mem == SGZ_MEM_HOST) SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	int sgz_spectrum_get_transform_result(sgz_spectrum* h, float* out, int mem)
	{
		if (!h || !out) return fail(SGZ_EINVAL, "null argument");
		SGZ_GUARD(h);
		const size_t want = size_t(h->tab.P) * h->tab.stateChannels;     // getTransformResult size / 2
		const size_t workStride = std::max<size_t>(h->tab.N, size_t(2) * h->tab.P);
		SGZ_CUDA(cudaMemcpy2DAsync(out, want * sizeof(float2), h->dWork, workStride * sizeof(float2), want * sizeof(float2), h->nStreams,
			mem == SGZ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
		if (mem == SGZ_MEM_HOST) SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	static int graphCopy(sgz_spectrum* h, int graph, float* out, const float* in, float2* base, int mem)
	{
		if (!h || (!out && !in)) return fail(SGZ_EINVAL, "null argument");
		if (graph < 0 || graph >= SGZ_LINE_END) return fail(SGZ_EINVAL, "graph %d out of range", graph);
		SGZ_GUARD(h);
		const size_t P = h->tab.P;
		const auto kind = out ? (mem == SGZ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost)
		                      : (mem == SGZ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
		if (out)
			SGZ_CUDA(cudaMemcpy2DAsync(out, P * sizeof(float2), base + size_t(graph) * P, 2 * P * sizeof(float2), P * sizeof(float2), h->nStreams, kind, h->stream));
		else
			SGZ_CUDA(cudaMemcpy2DAsync(base + size_t(graph) * P, 2 * P * sizeof(float2), in, P * sizeof(float2), P * sizeof(float2), h->nStreams, kind, h->stream));
		if (mem == SGZ_MEM_HOST) SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	int sgz_spectrum_get_results(sgz_spectrum* h, int graph, float* out, int mem) { return graphCopy(h, graph, out, nullptr, h ? h->dResults : nullptr, mem); }
	int sgz_spectrum_get_states(sgz_spectrum* h, int graph, float* out, int mem) { return graphCopy(h, graph, out, nullptr, h ? h->dStates : nullptr, mem); }
	int sgz_spectrum_set_states(sgz_spectrum* h, int graph, const float* in, int mem) { return graphCopy(h, graph, nullptr, in, h ? h->dStates : nullptr, mem); }

	int sgz_spectrum_clear_line_graph_states(sgz_spectrum* h)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		SGZ_GUARD(h);
		const size_t bytes = size_t(h->nStreams) * 2 * h->tab.P * sizeof(float2);
		SGZ_CUDA(cudaMemsetAsync(h->dStates, 0, bytes, h->stream));
		SGZ_CUDA(cudaMemsetAsync(h->dResults, 0, bytes, h->stream));
		return SGZ_OK;
	}

	int sgz_spectrum_clear_audio_state(sgz_spectrum* h)
	{
		int rc = sgz_spectrum_clear_line_graph_states(h);
		if (rc != SGZ_OK) return rc;
		SGZ_GUARD(h);
		const size_t workStride = std::max<size_t>(h->tab.N, size_t(2) * h->tab.P);
		SGZ_CUDA(cudaMemsetAsync(h->dCsf, 0, size_t(h->nStreams) * (h->tab.N + 1) * sizeof(float2), h->stream));
		SGZ_CUDA(cudaMemsetAsync(h->dWork, 0, size_t(h->nStreams) * workStride * sizeof(float2), h->stream));
		if (h->dResState) // cresonator.resetState(), TransformPair.h:183
			SGZ_CUDA(cudaMemsetAsync(h->dResState, 0, size_t(h->nStreams) * h->tab.stateChannels * h->tab.resonatorVectors * h->tab.P * sizeof(float2), h->stream));
		h->stage = 0;
		return SGZ_OK;
	}

	// launches resonatingDispatch over samples [first, first + count) of a device-resident buffer
	static int resonateBlock(sgz_spectrum* h, const float* dAudio, int layout, uint64_t nSamples, uint64_t streamStride, uint64_t first, uint32_t count)
	{
		if (count == 0) return SGZ_OK; // TransformDSP.inl:1246-1247
		const uint32_t half = h->tab.P * h->tab.resonatorVectors;
		const uint32_t total = half * h->tab.stateChannels;
		const uint32_t blocksX = (total + 255) / 256;
		int sms = 148;
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
		ResArgs a{};
		a.audio = dAudio; a.nSamples = nSamples; a.streamStride = streamStride; a.layout = layout;
		a.first = first; a.count = count;
		// Blocks of at least half a tile go to the tensor cores (k_resonate_tc): every time slice is a tile of kTcSpan samples
		// (the last one reaches back and reads zeros), k_resonate_combine chains them
		const uint32_t tcTiles = (h->dResTcA && h->resonatorPath != SGZ_RESONATOR_FP32 && count >= kTcSpan / 2) ? (count + kTcSpan - 1) / kTcSpan : 0;
		if (tcTiles > 0)
		{
			a.chunkLen = kTcSpan; a.chunks = tcTiles; a.chunk0 = tcTiles;
		}
		else
		{
			// enough time slices for ~2 waves of CTAs over the SMs, each at least 256 samples long
			uint64_t want = (uint64_t(sms) * 8 + uint64_t(blocksX) * h->nStreams - 1) / (uint64_t(blocksX) * h->nStreams);
			uint32_t chunks = uint32_t(std::max<uint64_t>(1, std::min<uint64_t>(want, count / 256)));
			if (chunks > 65535) chunks = 65535;
			a.chunks = chunks; a.chunkLen = (count + chunks - 1) / chunks;
			a.chunks = (count + a.chunkLen - 1) / a.chunkLen;
			a.chunk0 = 0;
		}
		if (a.chunks > 1 || tcTiles > 0)
		{
			const size_t need = size_t(h->nStreams) * a.chunks * total * sizeof(float2);
			if (h->resPartialBytes < need)
			{
				if (h->dResPartial) { SGZ_CUDA(cudaStreamSynchronize(h->stream)); cudaFree(h->dResPartial); h->dResPartial = nullptr; h->resPartialBytes = 0; }
				SGZ_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->dResPartial), need));
				h->resPartialBytes = need;
			}
		}
		if (tcTiles > 0)
		{
			ResTcArgs ta{};
			ta.audio = dAudio; ta.nSamples = nSamples; ta.streamStride = streamStride; ta.layout = layout; ta.first = first;
			ta.tiles = tcTiles; ta.count = count; ta.nStreams = h->nStreams; ta.aTiles = h->dResTcA; ta.powL = h->dResTcPowL;
			const uint64_t items = uint64_t((half + kTcRows - 1) / kTcRows) * h->nStreams * h->tab.stateChannels * tcTiles;
			const uint32_t ctas = uint32_t(std::min<uint64_t>(items, uint64_t(sms)));
			k_resonate_tc<<<ctas, kTcThreads, kTcSmemBytes, h->stream>>>(h->dev, ta, h->dResPartial);
		}
		const uint32_t fp32Chunks = a.chunks - a.chunk0; // slices the FP32 kernel computes
		if (fp32Chunks > 0)
		{
			if (h->tab.stateChannels == 2)
				k_resonate_pair<<<dim3((total / 2 + 255) / 256, fp32Chunks, h->nStreams), 256, 0, h->stream>>>(h->dev, a, h->dResState, h->dResPartial);
			else
				k_resonate<<<dim3(blocksX, fp32Chunks, h->nStreams), 256, 0, h->stream>>>(h->dev, a, h->dResState, h->dResPartial);
		}
		if (a.chunks > 1 || tcTiles > 0) k_resonate_combine<<<dim3(blocksX, h->nStreams), 256, 0, h->stream>>>(h->dev, a, h->dResState, h->dResPartial);
		SGZ_CUDA(cudaGetLastError());
		return SGZ_OK;
	}

	int sgz_spectrum_audio_entry_point(sgz_spectrum* h, const float* history, uint64_t n_history, uint64_t history_stride, const float* block, uint64_t n_block,
		uint64_t block_stride, int mem, int layout, float* frames, uint32_t max_frames, uint32_t* n_emitted)
	{
		if (!h || !block || !n_emitted || (max_frames && !frames)) return fail(SGZ_EINVAL, "null argument");
		*n_emitted = 0;
		const bool fft = h->cfg.algo == SGZ_ALGO_FFT, rsnt = h->cfg.algo == SGZ_ALGO_RSNT;
		if (fft && h->cfg.display_mode == SGZ_DISPLAY_COLOURSPECTRUM && !history) return fail(SGZ_EINVAL, "the FFT algorithm needs the audio-history view");
		const uint32_t P = h->tab.P;
		const uint64_t sampleBufferSize = h->cfg.sample_buffer_size;
		if (h->cfg.display_mode == SGZ_DISPLAY_COLOURSPECTRUM)
		{
			// TransformDSP.inl:1169-1204
			int64_t n = int64_t(n_block);
			uint64_t offset = 0;
			uint32_t emitted = 0;
			while (n > 0)
			{
				const int64_t numRemainingSamples = h->processedSamplesSinceLastFrame > sampleBufferSize ? 0 : int64_t(sampleBufferSize - h->processedSamplesSinceLastFrame);
				const int64_t availableSamples = numRemainingSamples + std::min<int64_t>(0, n - numRemainingSamples);
				int rc;
				if (rsnt && availableSamples > 0)
				{
					// resonatingDispatch(constant, { buffer[0] + offset, buffer[1] + offset }, availableSamples)
					SGZ_GUARD(h);
					const float* dBlock = nullptr;
					if ((rc = stageAudio(h, block, mem, n_block, block_stride, &dBlock, &h->dAudioStage, &h->audioStageBytes)) != SGZ_OK) return rc;
					if ((rc = resonateBlock(h, dBlock, layout, n_block, block_stride, offset, uint32_t(availableSamples))) != SGZ_OK) return rc;
				}
				h->processedSamplesSinceLastFrame += uint64_t(availableSamples);
				if (h->processedSamplesSinceLastFrame >= sampleBufferSize)
				{
					bool transformReady = true;
					if (fft)
					{
						// NOTE: `buffer`, not buffer + offset, exactly like the reference (:1192)
						// (the block keeps its own length as planar channel stride; only availableSamples of it count)
						rc = prepareCommon(h, history, n_history, history_stride, block, uint64_t(availableSamples), block_stride, mem, layout, true, n_block);
						if (rc < 0) return rc;
						transformReady = rc == SGZ_OK;
						if (transformReady && (rc = sgz_spectrum_transform(h)) != SGZ_OK) return rc;
					}
					if (transformReady)
					{
						// addAudioFrame: mapToLinearSpace + postProcessStdTransform + sfbuf.emplace_back(LineMain results)
						if ((rc = sgz_spectrum_map(h)) != SGZ_OK) return rc;
						if ((rc = sgz_spectrum_postprocess(h)) != SGZ_OK) return rc;
						if (emitted < max_frames)
						{
							SGZ_GUARD(h);
							SGZ_CUDA(cudaMemcpy2DAsync(frames + size_t(emitted) * P * 2, size_t(max_frames) * P * sizeof(float2), h->dResults, size_t(2) * P * sizeof(float2),
								size_t(P) * sizeof(float2), h->nStreams, cudaMemcpyDeviceToHost, h->stream));
						}
						++emitted;
					}
					h->processedSamplesSinceLastFrame = 0;
				}
				if (availableSamples <= 0 && numRemainingSamples <= 0 && sampleBufferSize == 0) break; // degenerate configuration
				offset += uint64_t(availableSamples);
				n -= availableSamples;
			}
			*n_emitted = emitted;
			SGZ_GUARD(h);
			SGZ_CUDA(cudaStreamSynchronize(h->stream));
			return SGZ_OK;
		}
		if (rsnt) return sgz_resonator_process(h, block, mem, layout, n_block, block_stride); // :1206-1209
		return SGZ_OK;
	}

	int sgz_spectrum_colour_columns(sgz_spectrum* h, const float* results, int mem, uint32_t n_frames, uint32_t pairs_per_view, const float* ratios,
		const float* colours, uint8_t* rgba, int out_mem)
	{
		if (!h || !results || !ratios || !colours || !rgba) return fail(SGZ_EINVAL, "null argument");
		if (pairs_per_view == 0 || h->nStreams % pairs_per_view != 0) return fail(SGZ_EINVAL, "n_streams must be a multiple of pairs_per_view");
		if (n_frames == 0) return SGZ_OK;
		if (n_frames > 65535 || h->nStreams / pairs_per_view > 65535) return fail(SGZ_EINVAL, "too many frames or views for one launch");
		SGZ_GUARD(h);
		const uint32_t P = h->tab.P, views = h->nStreams / pairs_per_view;
		const size_t resFloats = size_t(h->nStreams) * n_frames * 2 * P * 2;
		const size_t outBytes = size_t(views) * n_frames * P * 4;
		const size_t colBytes = size_t(pairs_per_view) * 18 * sizeof(float);
		const float* dRes = results;
		int rc;
		if (mem == SGZ_MEM_HOST)
		{
			if ((rc = ensureStage(h, &h->dResultStage, &h->resultStageBytes, resFloats * sizeof(float))) != SGZ_OK) return rc;
			SGZ_CUDA(cudaMemcpyAsync(h->dResultStage, results, resFloats * sizeof(float), cudaMemcpyHostToDevice, h->stream));
			dRes = h->dResultStage;
		}
		// colours + (for host callers) the output columns share the audio staging slot
		const size_t need = ((colBytes + 255) & ~size_t(255)) + (out_mem == SGZ_MEM_HOST ? outBytes : 0);
		if ((rc = ensureStage(h, &h->dAudioStage, &h->audioStageBytes, need)) != SGZ_OK) return rc;
		SGZ_CUDA(cudaMemcpyAsync(h->dAudioStage, colours, colBytes, cudaMemcpyHostToDevice, h->stream));
		uint8_t* dOut = out_mem == SGZ_MEM_HOST ? reinterpret_cast<uint8_t*>(h->dAudioStage) + ((colBytes + 255) & ~size_t(255)) : rgba;
		ColourArgs a{};
		a.results = reinterpret_cast<const float2*>(dRes); a.nFrames = n_frames; a.pairs = pairs_per_view; a.P = P;
		for (int k = 0; k < 6; ++k) a.ratios[k] = ratios[k];
		a.colours = h->dAudioStage; a.rgba = reinterpret_cast<uchar4*>(dOut);
		k_colour_columns<<<dim3((P + 255) / 256, n_frames, views), 256, 0, h->stream>>>(a);
		SGZ_CUDA(cudaGetLastError());
		if (out_mem == SGZ_MEM_HOST)
		{
			SGZ_CUDA(cudaMemcpyAsync(rgba, dOut, outBytes, cudaMemcpyDeviceToHost, h->stream));
			SGZ_CUDA(cudaStreamSynchronize(h->stream));
		}
		return SGZ_OK;
	}

	int sgz_resonator_process(sgz_spectrum* h, const float* audio, int mem, int layout, uint64_t n_samples, uint64_t stream_stride)
	{
		if (!h || !audio) return fail(SGZ_EINVAL, "null argument");
		if (h->cfg.algo != SGZ_ALGO_RSNT) return fail(SGZ_EUNSUPPORTED, "handle was not created with algo = SGZ_ALGO_RSNT");
		if (layout != SGZ_LAYOUT_PLANAR && layout != SGZ_LAYOUT_INTERLEAVED) return fail(SGZ_EINVAL, "bad layout");
		if (n_samples > 0xFFFFFFFFull) return fail(SGZ_EINVAL, "block too long");
		if (h->nStreams > 1 && stream_stride < 2 * n_samples) return fail(SGZ_EINVAL, "stream_stride smaller than one stream");
		if (layout == SGZ_LAYOUT_INTERLEAVED && ((stream_stride & 1) || (reinterpret_cast<uintptr_t>(audio) & 7)))
			return fail(SGZ_EINVAL, "interleaved audio must be 8-byte aligned with an even stream_stride");
		SGZ_GUARD(h);
		const float* dAudio = nullptr;
		int rc = stageAudio(h, audio, mem, n_samples, stream_stride, &dAudio, &h->dAudioStage, &h->audioStageBytes);
		if (rc != SGZ_OK) return rc;
		return resonateBlock(h, dAudio, layout, n_samples, stream_stride, 0, uint32_t(n_samples));
	}

	int sgz_resonator_set_path(sgz_spectrum* h, int path)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (path != SGZ_RESONATOR_AUTO && path != SGZ_RESONATOR_FP32) return fail(SGZ_EINVAL, "bad resonator path %d", path);
		h->resonatorPath = path;
		return SGZ_OK;
	}

	int sgz_resonator_get_state(sgz_spectrum* h, float* out, int mem)
	{
		if (!h || !out) return fail(SGZ_EINVAL, "null argument");
		if (!h->dResState) return fail(SGZ_EUNSUPPORTED, "handle was not created with algo = SGZ_ALGO_RSNT");
		SGZ_GUARD(h);
		const size_t bytes = size_t(h->nStreams) * h->tab.stateChannels * h->tab.resonatorVectors * h->tab.P * sizeof(float2);
		SGZ_CUDA(cudaMemcpyAsync(out, h->dResState, bytes, mem == SGZ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
		if (mem == SGZ_MEM_HOST) SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	// Host-resident callers: the streams are cut into chunks so that the upload of chunk c+1, the
	// kernel of chunk c and the download of chunk c-1 overlap (PCIe is full duplex); results are
	// identical to the single launch because streams are independent.
	// Launches the fused kernel for the streams [a.firstStream, a.firstStream + a.numStreams) on the handle's stream.
	// lineGraphs[k].results of the handle = the last frame (TransformPair.h:72-76): the rows of it that the output mode
	// kept in the caller's buffer are copied over; the others the kernel wrote there itself (resultRow)
	static int keepLastFrame(sgz_spectrum* h, const FusedArgs& a)
	{
		const size_t graph = size_t(h->tab.P) * sizeof(float2);
		const uint32_t F = a.nFrames, nS = a.numStreams;
		float2* dst = h->dResults + size_t(a.firstStream) * 2 * h->tab.P;
		if (a.outMode == SGZ_OUT_FULL)
			SGZ_CUDA(cudaMemcpy2DAsync(dst, 2 * graph, a.results + (size_t(a.firstStream) * F + (F - 1)) * 2 * h->tab.P, size_t(F) * 2 * graph, 2 * graph, nS, cudaMemcpyDeviceToDevice, h->stream));
		else if (a.outMode == SGZ_OUT_LINE_MAIN)
			SGZ_CUDA(cudaMemcpy2DAsync(dst, 2 * graph, a.results + (size_t(a.firstStream) * F + (F - 1)) * h->tab.P, size_t(F) * graph, graph, nS, cudaMemcpyDeviceToDevice, h->stream));
		else
			SGZ_CUDA(cudaMemcpyAsync(dst, a.results + size_t(a.firstStream) * 2 * h->tab.P, size_t(nS) * 2 * graph, cudaMemcpyDeviceToDevice, h->stream));
		return SGZ_OK;
	}

	// One k_fused8 launch over streams [a.firstStream, + a.numStreams) x a.nFrames frames.  The grid is the number of CTAs
	// the device holds at once (sgz_fused.cuh, FusedArgs): with at least that many streams the seams between CTA ranges
	// are chained inside the kernel; with fewer (a long recording rather than a wall of channels, or one chunk of the
	// host-buffer pipeline) parts start from zero and k_seam_carry / k_seam_apply carry the peak hold across afterwards.
	static int launchFusedFrames(sgz_spectrum* h, FusedArgs a)
	{
		const uint32_t P = h->tab.P, F = a.nFrames, nS = a.numStreams;
		const uint64_t items = uint64_t(nS) * F;
		const uint32_t capacity = std::max<uint32_t>(1, h->fusedCapacity);
		// Phase: .x is a peak hold like the others, .y a smoother (k_phase_smooth below, which needs every frame's filter input,
		// i.e. the full output)
		const bool phase = h->cfg.configuration == SGZ_CH_PHASE;
		FusedPlan plan{ nS, true, false };
#ifdef SGZ8_PER_STREAM_GRID
		if (false)
#else
		if (nS >= capacity)
#endif
		{
			plan.ctas = capacity;                 // balanced ranges of >= one stream's worth of frames each
			plan.cooperative = nS > capacity;     // some range ends mid-stream: CTAs wait on their neighbour's flag
#if SGZ8_DUAL
			// the headline shape (8192 points, Separate, Lanczos): two ranges per 1024-thread CTA
			plan.dual = h->fusedK1 == 8 && h->cfg.configuration == SGZ_CH_SEPARATE && h->cfg.bin_polation == SGZ_BIN_LANCZOS && capacity % 2 == 0 &&
				2 * ((Fused8GeomT<8>::smemBytes(P) + 127) & ~size_t(127)) <= size_t(227) * 1024;
#endif
		}
		else if (h->seamsOk && items / kFrameParallelMinRun > nS && !(phase && a.outMode != SGZ_OUT_FULL))
		{
			plan.ctas = uint32_t(std::min<uint64_t>(capacity, items / kFrameParallelMinRun));
			plan.chained = false;
			// a stream in at most a handful of parts, and not so long that decaying a state across it costs more than a launch
			// (not Phase: a part that is a whole stream would store its state itself, and Phase's parts keep the smoother's
			// input, not its state, in .y - k_seam_carry leaves .y alone)
			a.inlineCarry = (!phase && uint64_t(plan.ctas) <= uint64_t(4) * nS && F <= 512) ? 1u : 0u;
#if defined(SGZ_DIAG_SEAMS) && SGZ_DIAG_SEAMS == 1 // timing experiment: always k_seam_carry
			a.inlineCarry = 0;
#endif
		}
		a.lastResults = h->dResults; a.spare = h->dSpare;
		int rc = launchFused(h, a, &plan);
		if (rc != SGZ_OK) return rc;
		if (plan.chained) return keepLastFrame(h, a);
#if defined(SGZ_DIAG_SEAMS) && SGZ_DIAG_SEAMS == 2 // timing experiments: 2 = no carry at all (wrong results), 5 = k_seam_carry only (wrong results)
		return keepLastFrame(h, a);
#endif
		const int stereoFlag = (!phase && h->cfg.configuration != SGZ_CH_COMPLEX && h->tab.stateChannels == 2) ? 1 : 0; // the STEREO form of the hold
		if (!a.inlineCarry) k_seam_carry<<<dim3((P + 127) / 128, nS), 128, 0, h->stream>>>(h->dev, h->dSeam, h->dStates, F, nS, plan.ctas, stereoFlag, a.firstStream);
#if !defined(SGZ_DIAG_SEAMS) || SGZ_DIAG_SEAMS != 5
		{
			// a part is at most ceil(items / ctas) frames; slices of eight or more frames, sixteen at most (k_seam_apply)
			const uint32_t longest = uint32_t(std::min<uint64_t>(F, (items + plan.ctas - 1) / plan.ctas));
			const uint32_t slices = std::max<uint32_t>(1, std::min<uint32_t>(16, longest / 8));
			// (result rows are P float2 apart in every output mode: 16-byte accesses need an even P and a 16-byte aligned base)
			const bool wide = P % 2 == 0 && (reinterpret_cast<uintptr_t>(a.results) & 15) == 0;
			const bool full = a.outMode == SGZ_OUT_FULL;
			const dim3 grid((P / (wide ? 2 : 1) + 127) / 128, plan.ctas, slices);
			auto launch = [&](auto kern) { kern<<<grid, 128, 0, h->stream>>>(h->dev, a, plan.ctas); };
			if (wide && full) { if (stereoFlag) launch(k_seam_apply<2, true, true>); else launch(k_seam_apply<2, true, false>); }
			else if (wide) { if (stereoFlag) launch(k_seam_apply<2, false, true>); else launch(k_seam_apply<2, false, false>); }
			else if (stereoFlag) launch(k_seam_apply<1, false, true>);
			else launch(k_seam_apply<1, false, false>);
		}
#endif
		if (phase) k_phase_smooth<<<dim3((P + 127) / 128, nS), 128, 0, h->stream>>>(h->dev, a.results, h->dStates, F, a.firstStream);
		SGZ_CUDA(cudaGetLastError());
		return keepLastFrame(h, a);
	}

	// floats of `results` per stream in each output mode (include/sgz.h, SGZ_OUT_*)
	static size_t resultFloatsPerStream(int output, uint32_t n_frames, uint32_t P)
	{
		switch (output)
		{
		case SGZ_OUT_LINE_MAIN: return size_t(n_frames) * P * 2;
		case SGZ_OUT_LAST_FRAME: return size_t(2) * P * 2;
		default: return size_t(n_frames) * 2 * P * 2;
		}
	}

	// waits for whatever a failed call left in flight on the handle's streams (copies into or out of caller memory)
	static int drained(sgz_spectrum* h, int rc)
	{
		if (h->copyIn) cudaStreamSynchronize(h->copyIn);
		if (h->stream) cudaStreamSynchronize(h->stream);
		if (h->copyOut) cudaStreamSynchronize(h->copyOut);
		return rc;
	}

	static int processFramesPipelined(sgz_spectrum* h, const float* audio, int layout, uint64_t n_samples, uint64_t stream_stride, uint32_t hop,
		uint32_t n_frames, float* results, int output)
	{
		const uint32_t P = h->tab.P, S = h->nStreams;
		const size_t audioFloats = size_t(stream_stride) * (S - 1) + size_t(n_samples) * 2;
		const size_t resPerStream = resultFloatsPerStream(output, n_frames, P);
		int rc = ensureStage(h, &h->dAudioStage, &h->audioStageBytes, audioFloats * sizeof(float));
		if (rc != SGZ_OK) return rc;
		if ((rc = ensureStage(h, &h->dResultStage, &h->resultStageBytes, resPerStream * S * sizeof(float))) != SGZ_OK) return rc;
		if (!h->copyIn) SGZ_CUDA(cudaStreamCreateWithFlags(&h->copyIn, cudaStreamNonBlocking));
		if (!h->copyOut) SGZ_CUDA(cudaStreamCreateWithFlags(&h->copyOut, cudaStreamNonBlocking));
		// chunks of streams flow through copy-in -> kernels -> copy-out; the first copy-in and the last copy-out are not
		// overlapped, so smaller chunks shorten that fill / drain.  A chunk of a few dozen streams has fewer streams than
		// the GPU holds CTAs, which is what launchFusedFrames' unchained mode is for
		const uint32_t chunks = std::max<uint32_t>(1, std::min<uint32_t>(h->pipeChunks, S / 8));
		while (h->events.size() < 2 * chunks + 1)
		{
			cudaEvent_t e;
			SGZ_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
			h->events.push_back(e);
		}
		// the copy streams start after whatever the caller queued on the handle's stream
		SGZ_CUDA(cudaEventRecord(h->events[2 * chunks], h->stream));
		SGZ_CUDA(cudaStreamWaitEvent(h->copyIn, h->events[2 * chunks], 0));
		auto fail2 = [&](cudaError_t e, const char* what) { return drained(h, fail(SGZ_ECUDA, "%s failed: %s", what, cudaGetErrorString(e))); };
		for (uint32_t c = 0; c < chunks; ++c)
		{
			const uint32_t s0 = uint32_t(uint64_t(S) * c / chunks), s1 = uint32_t(uint64_t(S) * (c + 1) / chunks);
			if (s1 == s0) continue;
			const size_t off = size_t(s0) * stream_stride;
			const size_t floats = size_t(stream_stride) * (s1 - s0 - 1) + size_t(n_samples) * 2;
			cudaError_t e;
			if ((e = cudaMemcpyAsync(h->dAudioStage + off, audio + off, floats * sizeof(float), cudaMemcpyHostToDevice, h->copyIn)) != cudaSuccess) return fail2(e, "copy-in");
			if ((e = cudaEventRecord(h->events[2 * c], h->copyIn)) != cudaSuccess) return fail2(e, "event record");
			if ((e = cudaStreamWaitEvent(h->stream, h->events[2 * c], 0)) != cudaSuccess) return fail2(e, "stream wait");
			FusedArgs a{};
			a.audio = h->dAudioStage; a.nSamples = n_samples; a.streamStride = stream_stride; a.layout = layout;
			a.hop = hop; a.nFrames = n_frames; a.states = h->dStates; a.results = reinterpret_cast<float2*>(h->dResultStage);
			a.outMode = uint32_t(output);
			a.firstStream = s0; a.numStreams = s1 - s0;
			if ((rc = launchFusedFrames(h, a)) != SGZ_OK) return drained(h, rc);
			if ((e = cudaEventRecord(h->events[2 * c + 1], h->stream)) != cudaSuccess) return fail2(e, "event record");
			if ((e = cudaStreamWaitEvent(h->copyOut, h->events[2 * c + 1], 0)) != cudaSuccess) return fail2(e, "stream wait");
			if ((e = cudaMemcpyAsync(results + resPerStream * s0, h->dResultStage + resPerStream * s0, resPerStream * (s1 - s0) * sizeof(float),
				cudaMemcpyDeviceToHost, h->copyOut)) != cudaSuccess) return fail2(e, "copy-out");
		}
		SGZ_CUDA(cudaStreamSynchronize(h->copyOut));
		SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	// one frame's lineGraphs results (h->dResults, [S][2][P]) -> its place in `dRes` for the staged / resonator paths
	static int storeStagedFrame(sgz_spectrum* h, float* dRes, int output, uint32_t f, uint32_t n_frames)
	{
		const uint32_t P = h->tab.P;
		const size_t graph = size_t(P) * sizeof(float2);
		if (output == SGZ_OUT_FULL)
			SGZ_CUDA(cudaMemcpy2DAsync(dRes + size_t(f) * 2 * P * 2, size_t(n_frames) * 2 * graph, h->dResults, 2 * graph, 2 * graph, h->nStreams, cudaMemcpyDeviceToDevice, h->stream));
		else if (output == SGZ_OUT_LINE_MAIN)
			SGZ_CUDA(cudaMemcpy2DAsync(dRes + size_t(f) * P * 2, size_t(n_frames) * graph, h->dResults, 2 * graph, graph, h->nStreams, cudaMemcpyDeviceToDevice, h->stream));
		else if (f + 1 == n_frames)
			SGZ_CUDA(cudaMemcpyAsync(dRes, h->dResults, size_t(h->nStreams) * 2 * graph, cudaMemcpyDeviceToDevice, h->stream));
		return SGZ_OK;
	}

	int sgz_spectrum_process_frames_select(sgz_spectrum* h, const float* audio, int mem, int layout, uint64_t n_samples, uint64_t stream_stride, uint32_t hop,
		uint32_t n_frames, float* results, int results_mem, int output)
	{
		if (!h || !audio || !results) return fail(SGZ_EINVAL, "null argument");
		if (layout != SGZ_LAYOUT_PLANAR && layout != SGZ_LAYOUT_INTERLEAVED) return fail(SGZ_EINVAL, "bad layout");
		if (output != SGZ_OUT_FULL && output != SGZ_OUT_LINE_MAIN && output != SGZ_OUT_LAST_FRAME) return fail(SGZ_EINVAL, "bad output selector %d", output);
		if (n_frames == 0) return SGZ_OK;
		const uint32_t W = h->tab.W, P = h->tab.P;
		const bool rsnt = h->cfg.algo == SGZ_ALGO_RSNT;
		if (rsnt && n_samples < uint64_t(n_frames) * hop) return fail(SGZ_EINVAL, "n_samples %llu < n_frames*hop", (unsigned long long)n_samples);
		if (!rsnt && n_samples < uint64_t(n_frames - 1) * hop + W) return fail(SGZ_EINVAL, "n_samples %llu < (n_frames-1)*hop + window_size", (unsigned long long)n_samples);
		if (h->nStreams > 1 && stream_stride < 2 * n_samples) return fail(SGZ_EINVAL, "stream_stride smaller than one stream");
		if (layout == SGZ_LAYOUT_INTERLEAVED && ((stream_stride & 1) || (reinterpret_cast<uintptr_t>(audio) & 7)))
			return fail(SGZ_EINVAL, "interleaved audio must be 8-byte aligned with an even stream_stride");
		SGZ_GUARD(h);
		if (!rsnt && h->fusedK1 != 0 && mem == SGZ_MEM_HOST && results_mem == SGZ_MEM_HOST && h->nStreams >= 8)
			return processFramesPipelined(h, audio, layout, n_samples, stream_stride, hop, n_frames, results, output);
		const float* dAudio = nullptr;
		int rc = stageAudio(h, audio, mem, n_samples, stream_stride, &dAudio, &h->dAudioStage, &h->audioStageBytes);
		if (rc != SGZ_OK) return rc;
		const size_t resFloats = size_t(h->nStreams) * resultFloatsPerStream(output, n_frames, P);
		float* dRes = results;
		if (results_mem == SGZ_MEM_HOST)
		{
			rc = ensureStage(h, &h->dResultStage, &h->resultStageBytes, resFloats * sizeof(float));
			if (rc != SGZ_OK) return drained(h, rc);
			dRes = h->dResultStage;
		}
		if (rsnt)
		{
			// audioEntryPoint, RSNT (TransformDSP.inl:1172-1204): resonate one blob of `hop` samples, then
			// addAudioFrame = snapshot + post-process
			for (uint32_t f = 0; f < n_frames; ++f)
			{
				if ((rc = resonateBlock(h, dAudio, layout, n_samples, stream_stride, uint64_t(f) * hop, hop)) != SGZ_OK) return drained(h, rc);
				if ((rc = sgz_spectrum_map(h)) != SGZ_OK) return drained(h, rc);
				if ((rc = sgz_spectrum_postprocess(h)) != SGZ_OK) return drained(h, rc);
				if ((rc = storeStagedFrame(h, dRes, output, f, n_frames)) != SGZ_OK) return drained(h, rc);
			}
		}
		else if (h->fusedK1 != 0)
		{
			FusedArgs a{};
			a.audio = dAudio; a.nSamples = n_samples; a.streamStride = stream_stride; a.layout = layout;
			a.hop = hop; a.nFrames = n_frames; a.states = h->dStates; a.results = reinterpret_cast<float2*>(dRes);
			a.outMode = uint32_t(output);
			a.firstStream = 0; a.numStreams = h->nStreams;
			if ((rc = launchFusedFrames(h, a)) != SGZ_OK) return drained(h, rc);
		}
		else
		{
			// staged fallback ON THE GPU (every configuration / size): four kernels per frame
			for (uint32_t f = 0; f < n_frames; ++f)
			{
				PrepareArgs a{};
				a.layout = layout; a.hist = dAudio; a.nHist = n_samples; a.histStride = stream_stride;
				a.histSkip = uint32_t(uint64_t(f) * hop); a.histCount = W; a.prelimCount = 0;
				if (uint64_t(f) * hop > 0xFFFFFFFFull) return drained(h, fail(SGZ_EUNSUPPORTED, "frame offset exceeds 32 bits"));
				const dim3 block(256), grid((h->tab.N + 1 + 255) / 256, h->nStreams);
				k_prepare<<<grid, block, 0, h->stream>>>(h->dev, a, h->dCsf);
				h->stage = 1;
				if ((rc = sgz_spectrum_transform(h)) != SGZ_OK) return drained(h, rc);
				if ((rc = sgz_spectrum_map(h)) != SGZ_OK) return drained(h, rc);
				if ((rc = sgz_spectrum_postprocess(h)) != SGZ_OK) return drained(h, rc);
				if ((rc = storeStagedFrame(h, dRes, output, f, n_frames)) != SGZ_OK) return drained(h, rc);
			}
		}
		SGZ_CUDA(cudaGetLastError());
		if (results_mem == SGZ_MEM_HOST)
		{
			SGZ_CUDA(cudaMemcpyAsync(results, dRes, resFloats * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
			SGZ_CUDA(cudaStreamSynchronize(h->stream));
		}
		return SGZ_OK;
	}

	int sgz_spectrum_reserve(sgz_spectrum* h, uint64_t n_samples, uint64_t stream_stride, uint32_t n_frames)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (h->nStreams > 1 && stream_stride < 2 * n_samples) return fail(SGZ_EINVAL, "stream_stride smaller than one stream");
		SGZ_GUARD(h);
		const uint32_t P = h->tab.P;
		const size_t audioFloats = size_t(stream_stride) * (h->nStreams - 1) + size_t(n_samples) * 2;
		int rc = ensureStage(h, &h->dAudioStage, &h->audioStageBytes, audioFloats * sizeof(float));
		if (rc != SGZ_OK) return rc;
		if ((rc = ensureStage(h, &h->dResultStage, &h->resultStageBytes, size_t(h->nStreams) * resultFloatsPerStream(SGZ_OUT_FULL, n_frames, P) * sizeof(float))) != SGZ_OK) return rc;
		if (!h->copyIn) SGZ_CUDA(cudaStreamCreateWithFlags(&h->copyIn, cudaStreamNonBlocking));
		if (!h->copyOut) SGZ_CUDA(cudaStreamCreateWithFlags(&h->copyOut, cudaStreamNonBlocking));
		while (h->events.size() < 2 * h->pipeChunks + 1)
		{
			cudaEvent_t e;
			SGZ_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
			h->events.push_back(e);
		}
		if (h->cfg.algo == SGZ_ALGO_RSNT)
		{
			// time-slice scratch of the resonator for blocks of up to n_samples: tiles of the tensor-core kernel, or the
			// FP32 kernels' slices of at least 256 samples
			const size_t total = size_t(P) * h->tab.resonatorVectors * h->tab.stateChannels;
			const size_t slices = std::max<size_t>((n_samples + kTcSpan - 1) / kTcSpan, std::min<size_t>(65535, n_samples / 256 + 1));
			const size_t need = size_t(h->nStreams) * slices * total * sizeof(float2);
			if (h->resPartialBytes < need)
			{
				if (h->dResPartial) { SGZ_CUDA(cudaStreamSynchronize(h->stream)); cudaFree(h->dResPartial); h->dResPartial = nullptr; h->resPartialBytes = 0; }
				SGZ_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->dResPartial), need));
				h->resPartialBytes = need;
			}
		}
		return SGZ_OK;
	}

	int sgz_spectrum_process_frames(sgz_spectrum* h, const float* audio, int mem, int layout, uint64_t n_samples, uint64_t stream_stride, uint32_t hop,
		uint32_t n_frames, float* results, int results_mem)
	{
		return sgz_spectrum_process_frames_select(h, audio, mem, layout, n_samples, stream_stride, hop, n_frames, results, results_mem, SGZ_OUT_FULL);
	}

	// ---- device-resident audio history ---------------------------------------------------------
	int sgz_spectrum_history_reserve(sgz_spectrum* h, uint64_t capacity)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (capacity < h->tab.W) return fail(SGZ_EINVAL, "history capacity %llu is smaller than the window (%u)", (unsigned long long)capacity, h->tab.W);
		if (capacity > (1ull << 28)) return fail(SGZ_EINVAL, "history capacity %llu exceeds 2^28 samples", (unsigned long long)capacity);
		SGZ_GUARD(h);
		const size_t bytes = size_t(h->nStreams) * 2 * capacity * sizeof(float2);
		if (h->ringCapacity != capacity)
		{
			if (h->dRing) { SGZ_CUDA(cudaStreamSynchronize(h->stream)); cudaFree(h->dRing); h->dRing = nullptr; h->ringCapacity = 0; }
			SGZ_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->dRing), bytes));
			h->ringCapacity = capacity;
		}
		SGZ_CUDA(cudaMemsetAsync(h->dRing, 0, bytes, h->stream)); // a fresh history is silence
		h->ringTotal = 0;
		h->ringLastFrame = 0;
		return SGZ_OK;
	}

	int sgz_spectrum_history_append(sgz_spectrum* h, const float* block, int mem, int layout, uint64_t n_block, uint64_t stream_stride)
	{
		if (!h || !block) return fail(SGZ_EINVAL, "null argument");
		if (!h->dRing) return fail(SGZ_ESTATE, "history_append before history_reserve");
		if (layout != SGZ_LAYOUT_PLANAR && layout != SGZ_LAYOUT_INTERLEAVED) return fail(SGZ_EINVAL, "bad layout");
		if (n_block == 0) return SGZ_OK;
		if (h->nStreams > 1 && stream_stride < 2 * n_block) return fail(SGZ_EINVAL, "stream_stride smaller than one stream");
		if (layout == SGZ_LAYOUT_INTERLEAVED && ((stream_stride & 1) || (reinterpret_cast<uintptr_t>(block) & 7)))
			return fail(SGZ_EINVAL, "interleaved audio must be 8-byte aligned with an even stream_stride");
		SGZ_GUARD(h);
		const float* dBlock = nullptr;
		int rc = stageAudio(h, block, mem, n_block, stream_stride, &dBlock, &h->dAudioStage, &h->audioStageBytes);
		if (rc != SGZ_OK) return rc;
		const uint64_t C = h->ringCapacity;
		const uint64_t skip = n_block > C ? n_block - C : 0; // only the last C samples of an oversized block survive
		const uint64_t n = n_block - skip;
		const dim3 grid(unsigned((n + 255) / 256), h->nStreams);
		k_history_append<<<grid, 256, 0, h->stream>>>(h->dRing, C, h->ringTotal % C, dBlock, layout, n_block, stream_stride, skip);
		SGZ_CUDA(cudaGetLastError());
		if (mem == SGZ_MEM_HOST) SGZ_CUDA(cudaStreamSynchronize(h->stream)); // the caller may reuse its buffer on return
		h->ringTotal += n_block;
		return SGZ_OK;
	}

	int sgz_spectrum_history_info(sgz_spectrum* h, uint64_t* capacity, uint64_t* total_appended, uint64_t* last_frame_position)
	{
		if (!h) return fail(SGZ_EINVAL, "null handle");
		if (capacity) *capacity = h->ringCapacity;
		if (total_appended) *total_appended = h->ringTotal;
		if (last_frame_position) *last_frame_position = h->ringLastFrame;
		return SGZ_OK;
	}

	int sgz_spectrum_history_view(sgz_spectrum* h, uint64_t n_last, const float** device_audio, uint64_t* stream_stride)
	{
		if (!h || !device_audio || !stream_stride) return fail(SGZ_EINVAL, "null argument");
		if (!h->dRing) return fail(SGZ_ESTATE, "history_view before history_reserve");
		const uint64_t C = h->ringCapacity;
		if (n_last > C) return fail(SGZ_EINVAL, "history holds %llu samples, %llu requested", (unsigned long long)C, (unsigned long long)n_last);
		// the run ends at the write head; before the first wrap the samples ahead of time zero are the silence
		// the ring was cleared to
		const uint64_t head = h->ringTotal % C;
		const uint64_t start = (head + C - n_last) % C;
		*device_audio = reinterpret_cast<const float*>(h->dRing + start);
		*stream_stride = 4 * C; // floats between streams: 2 C stereo samples
		return SGZ_OK;
	}

	int sgz_spectrum_history_read(sgz_spectrum* h, uint64_t n_last, float* out, int mem)
	{
		if (!out) return fail(SGZ_EINVAL, "null argument");
		const float* view = nullptr; uint64_t stride = 0;
		int rc = sgz_spectrum_history_view(h, n_last, &view, &stride);
		if (rc != SGZ_OK) return rc;
		if (n_last == 0) return SGZ_OK;
		SGZ_GUARD(h);
		SGZ_CUDA(cudaMemcpy2DAsync(out, size_t(n_last) * sizeof(float2), view, size_t(stride) * sizeof(float), size_t(n_last) * sizeof(float2), h->nStreams,
			mem == SGZ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
		if (mem == SGZ_MEM_HOST) SGZ_CUDA(cudaStreamSynchronize(h->stream));
		return SGZ_OK;
	}

	int sgz_spectrum_history_process(sgz_spectrum* h, uint32_t hop, uint32_t max_frames, float* results, int results_mem, uint32_t* n_frames)
	{
		if (!h || !n_frames || (max_frames && !results)) return fail(SGZ_EINVAL, "null argument");
		*n_frames = 0;
		if (!h->dRing) return fail(SGZ_ESTATE, "history_process before history_reserve");
		if (hop == 0) return fail(SGZ_EINVAL, "hop must be positive");
		if (h->cfg.algo != SGZ_ALGO_FFT) return fail(SGZ_EUNSUPPORTED, "history_process is the FFT path; the resonator consumes blocks directly (sgz_spectrum_audio_entry_point)");
		const uint64_t C = h->ringCapacity, W = h->tab.W, T = h->ringTotal;
		// frames fall due every `hop` samples after the last one (the processedSamplesSinceLastFrame counter of
		// audioEntryPoint, TransformDSP.inl:1172-1204); the frame at boundary b transforms samples [b - W, b)
		const uint64_t pending = (T - h->ringLastFrame) / hop;
		const uint64_t n = std::min<uint64_t>(pending, max_frames);
		if (n == 0) return SGZ_OK;
		const uint64_t firstBoundary = h->ringLastFrame + hop;
		const uint64_t span = (n - 1) * hop + W;
		// every sample of [firstBoundary - W, T) must still be in the ring (times before zero are the silence
		// the ring was cleared to, and need their slots unwritten just the same)
		if (T + W - firstBoundary > C)
			return fail(SGZ_ESTATE, "history overrun: the window of the oldest pending frame was overwritten (capacity %llu, %llu samples behind); "
				"process more often or reserve a longer history", (unsigned long long)C, (unsigned long long)(T + W - firstBoundary));
		// ring position of absolute time firstBoundary - W (may be negative in absolute terms: add multiples of C)
		const uint64_t start = (firstBoundary % C + C - W % C) % C;
		const float* audio = reinterpret_cast<const float*>(h->dRing + start);
		// start + span <= 2 C: span <= C by the overrun check
		int rc = sgz_spectrum_process_frames(h, audio, SGZ_MEM_DEVICE, SGZ_LAYOUT_INTERLEAVED, span, 4 * C, hop, uint32_t(n), results, results_mem);
		if (rc != SGZ_OK) return rc;
		h->ringLastFrame += n * hop;
		*n_frames = uint32_t(n);
		return SGZ_OK;
	}

	// ---- frequency tracker ---------------------------------------------------------------------
	int sgz_spectrum_track_peaks(sgz_spectrum* h, int graph, const double* mouse_fraction, double smoothing_ms, double fps, sgz_tracker_state* states,
		sgz_tracker_result* results)
	{
		if (!h || !mouse_fraction || !results) return fail(SGZ_EINVAL, "null argument");
		if (graph < SGZ_TRACK_LINE_MAIN || graph > SGZ_TRACK_TRANSFORM) return fail(SGZ_EINVAL, "bad graph");
		const Tables& t = h->tab;
		const bool fft = h->cfg.algo == SGZ_ALGO_FFT;
		const bool raw = h->cfg.configuration != SGZ_CH_COMPLEX && fft && graph == SGZ_TRACK_TRANSFORM;
		if (raw && h->stage < 2) return fail(SGZ_ESTATE, "the raw-FFT tracker reads audioMemory: call sgz_spectrum_transform first");
		if (states && !(fps > 0)) return fail(SGZ_EINVAL, "fps must be positive");
		const long long P = t.P, N = t.N;
		if (P < 2) return fail(SGZ_EINVAL, "the tracker needs at least two display points");
		SGZ_GUARD(h);
		const uint32_t S = h->nStreams;
		const size_t boundsBytes = size_t(S) * sizeof(TrackBounds), recBytes = size_t(S) * sizeof(TrackRecord);
		if (!h->dTrack) SGZ_CUDA(cudaMalloc(&h->dTrack, boundsBytes + recBytes));
		TrackBounds* dBounds = static_cast<TrackBounds*>(h->dTrack);
		TrackRecord* dRec = reinterpret_cast<TrackRecord*>(static_cast<char*>(h->dTrack) + boundsBytes);
		const double sampleRate = double(h->cfg.sample_rate);
		auto roundHalfUp = [](double x) { return (long long)std::floor(x + 0.5); };
		auto clampTo = [](long long x, long long lo, long long hi) { return x < lo ? lo : (x > hi ? hi : x); };
		std::vector<TrackBounds> bounds(S);
		std::vector<double> mouse(S);
		for (uint32_t s = 0; s < S; ++s)
		{
			const double mf = std::min(1.0, std::max(0.0, mouse_fraction[s]));
			mouse[s] = mf;
			TrackBounds b{};
			if (!raw)
			{
				// window of +-3 % of the axis around the mouse, in display points
				const unsigned long long pivot = (unsigned long long)roundHalfUp(double(P) * mf), range = (unsigned long long)roundHalfUp(double(P) * 0.03);
				b.lo = range > pivot ? 0 : (long long)(pivot - range);
				b.hiEx = range + pivot > (unsigned long long)P ? P : (long long)(range + pivot);
				b.rightBorder = b.hiEx - 1;
				b.end = P;
				if (b.lo >= b.hiEx) return fail(SGZ_EINVAL, "empty tracker window (axis_points too small)");
			}
			else
			{
				// the same window expressed in FFT bins through the display-point frequencies
				long long lo = roundHalfUp(double(P) * (mf - 0.03)), hi = roundHalfUp(double(P) * (mf + 0.03));
				lo = roundHalfUp(double(N) * t.mappedFreq[size_t(clampTo(lo, 0, P - 1))] / sampleRate);
				hi = roundHalfUp(double(N) * t.mappedFreq[size_t(clampTo(hi, 0, P - 1))] / sampleRate);
				lo = clampTo(lo, 0, N); hi = clampTo(hi, 0, N);
				b.lo = lo; b.hiEx = hi + 1; b.rightBorder = hi - 1; b.end = N;
				if (b.lo >= b.hiEx) return fail(SGZ_EINVAL, "empty tracker window");
			}
			bounds[s] = b;
		}
		SGZ_CUDA(cudaMemcpyAsync(dBounds, bounds.data(), boundsBytes, cudaMemcpyHostToDevice, h->stream));
		const int shownGraph = graph == SGZ_TRACK_TRANSFORM ? SGZ_TRACK_LINE_MAIN : graph;
		const dim3 block(128), grid((S * 32 + 127) / 128);
		if (raw) k_track_peaks<true><<<grid, block, 0, h->stream>>>(h->dCsf, uint64_t(N) + 1, dBounds, dRec, S);
		else k_track_peaks<false><<<grid, block, 0, h->stream>>>(h->dResults + size_t(shownGraph) * P, uint64_t(2) * P, dBounds, dRec, S);
		SGZ_CUDA(cudaGetLastError());
		std::vector<TrackRecord> rec(S);
		SGZ_CUDA(cudaMemcpyAsync(rec.data(), dRec, recBytes, cudaMemcpyDeviceToHost, h->stream));
		SGZ_CUDA(cudaStreamSynchronize(h->stream));

		const double nyquist = sampleRate * 0.5;
		for (uint32_t s = 0; s < S; ++s)
		{
			const TrackRecord& r = rec[s];
			sgz_tracker_result o{};
			o.peak_offset = r.peak;
			o.searched_raw_fft = raw ? 1 : 0;
			double freq, dbs;
			if (!raw)
			{
				const size_t k = size_t(r.peak);
				freq = t.mappedFreq[k];
				const bool last = r.peak == P - 1;
				const float width = t.mappedFreq[last ? k : k + 1] - t.mappedFreq[last ? k - 1 : k]; // float difference, as the reference
				o.peak_deviance = width;
				if (fft && h->cfg.bin_polation != SGZ_BIN_LANCZOS) o.peak_deviance = std::max(o.peak_deviance, 0.5 * double(t.N) / double(P));
				o.peak_x = double(r.peak);
				o.peak_slope = t.slope[k];
				dbs = double(r.at.x) * (h->cfg.high_dbs - h->cfg.low_dbs) + h->cfg.low_dbs; // results are fractions of the dB range
			}
			else
			{
				const float invSize = static_cast<float>(double(t.windowKernelScale) / (double(t.W) * 0.5));
				auto levelDb = [invSize](float2 z) -> float { return 20 * std::log10(std::abs(std::complex<float>(z.x, z.y) * invSize)); };
				const float a = levelDb(r.prev), bb = levelDb(r.at), c = levelDb(r.next);
				const double phi = 0.5 * (a - c) / (a - 2 * bb + c); // vertex of the parabola through the three log magnitudes
				o.peak_fraction = 2 * (double(r.peak) + (std::isnormal(phi) ? phi : 0)) / double(N);
				freq = 0.5 * o.peak_fraction * sampleRate;
				dbs = bb - 0.25 * (a - c) * phi;
				if (!std::isnormal(dbs)) dbs = 20 * std::log10(std::abs(std::complex<float>(r.at.x, r.at.y)) / (double(N) * 0.5));
				// display coordinate of that frequency: inverse of the axis mapping (TransformConstant.h:125-182)
				double arg = o.peak_fraction;
				if (h->cfg.view_scaling == SGZ_VIEW_LOGARITHMIC) arg = std::log(o.peak_fraction * nyquist / h->cfg.min_freq) / std::log(nyquist / h->cfg.min_freq);
				o.peak_x = (arg - h->cfg.view_left) / (h->cfg.view_right - h->cfg.view_left) * double(P - 1);
				o.peak_slope = t.slope[size_t(clampTo(roundHalfUp(o.peak_x), 0, P - 1))];
				dbs += 20 * std::log10(o.peak_slope);
				o.peak_deviance = 2 * 0.01 * ((1.0 - o.peak_fraction) / double(N)) * sampleRate + 0.001;
			}
			o.peak_is_complex = freq > nyquist ? 1 : 0;
			if (o.peak_is_complex) freq = sampleRate - freq;
			if (dbs > 1000) dbs = INFINITY; else if (dbs < -1000) dbs = -INFINITY;
			o.peak_frequency = freq; o.peak_dbs = dbs;
			o.smoothed_frequency = freq; o.smoothed_dbs = dbs;
			if (states)
			{
				sgz_tracker_state& st = states[s];
				if (!std::isnormal(st.current_peak)) st.current_peak = 0;
				if (!std::isnormal(st.frequency_state[7])) std::fill(st.frequency_state, st.frequency_state + 8, 0.0);
				if (!std::isnormal(st.peak_dbs_state[7])) std::fill(st.peak_dbs_state, st.peak_dbs_state + 8, 0.0);
				const double holdPole = std::exp(-1.0 / (smoothing_ms * 10 * 0.001 * fps)), smoothPole = std::exp(-1.0 / (smoothing_ms / 5 * 0.001 * fps));
				const double linear = std::pow(10, dbs / 20);
				st.current_peak *= holdPole;
				if (linear > st.current_peak) { st.current_peak = 1.2 * (linear / holdPole); st.current_frequency = freq; st.current_peak_dbs = dbs; }
				double in = st.current_frequency;
				for (double& x : st.frequency_state) { x = in + smoothPole * (x - in); in = x; }
				in = st.current_peak_dbs;
				for (double& x : st.peak_dbs_state) { x = in + smoothPole * (x - in); in = x; }
				o.smoothed_frequency = st.frequency_state[7];
				o.smoothed_dbs = st.peak_dbs_state[7];
			}
			results[s] = o;
		}
		return SGZ_OK;
	}
}
