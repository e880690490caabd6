// sgz_noalloc.cu — a forward FFT entry point for callers that must not allocate device memory.
//
// Who needs this: code that runs where cudaMalloc / cudaFree are off limits — an audio or render
// callback with a hard deadline (cudaMalloc takes a driver lock and may synchronise the whole
// device), a plugin host that hands the library pre-allocated device buffers and forbids any other
// allocation, or a process whose memory pool was sized up front and is now full.  The handle-based
// API (sgz_spectrum_create) allocates its working buffers with cudaMalloc; this entry point needs
// no handle and allocates nothing: its ping-pong buffer is a fixed array that lives inside the
// library image on every GPU that loads it.
//
// It is the handle-free twin of sgz_spectrum_transform's large-size path (Stockham radix-2 passes
// through global memory).  There is no counterpart in the reference.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>

#include "../../include/sgz.h"
#include "sgz_device.cuh"

namespace sgz
{
	int fail(int code, const char* fmt, ...); // sgz_spectrum.cu

	// The whole scratch: SGZ_NOALLOC_MAX_POINTS complex floats = 8 MiB per GPU, held until the
	// process exits.  In host code this name is only a stand-in for the symbol; the device address
	// comes from cudaGetSymbolAddress below.
	static __device__ __align__(256) float2 g_noallocScratch[SGZ_NOALLOC_MAX_POINTS];

	// One Stockham radix-2 pass (same butterfly as k_fft_global_pass in sgz_spectrum.cu) with the
	// twiddle computed in place of a table lookup, so no table has to live anywhere.
	__global__ void k_fft_pass_noalloc(const float2* __restrict__ src, float2* __restrict__ dst, uint32_t N, uint32_t n, uint32_t s)
	{
		const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; // butterfly p * s + q
		if (b >= (N >> 1)) return;
		const float2* x = src + uint64_t(blockIdx.y) * N;
		float2* y = dst + uint64_t(blockIdx.y) * N;
		const uint32_t q = b % s, p = b / s;
		const uint32_t n1 = n >> 1;
		float sn, cs;
		sincospif(-2.0f * float(p) / float(n), &sn, &cs); // W_n^p; p / n is exact in float (n <= 2^20)
		const float2 u = x[q + s * p], v = x[q + s * (p + n1)];
		y[q + s * (2 * p)] = cadd(u, v);
		y[q + s * (2 * p + 1)] = cmul(csub(u, v), make_float2(cs, sn));
	}

	// per-device address of g_noallocScratch, looked up once per device (host threads may race to the first call on
	// a device: they all store the same value)
	static std::atomic<float2*> g_scratchAddress[64];
}

extern "C" int sgz_fft_forward_noalloc(const float* d_in, float* d_out, uint32_t n, uint32_t batch, void* cuda_stream)
{
	using namespace sgz;
	if (!d_in || !d_out) return fail(SGZ_EINVAL, "sgz_fft_forward_noalloc: null buffer");
	if (d_in == d_out) return fail(SGZ_EINVAL, "sgz_fft_forward_noalloc: in-place transforms are not supported, pass distinct buffers");
	if (n < 2 || (n & (n - 1)) != 0) return fail(SGZ_EINVAL, "sgz_fft_forward_noalloc: n = %u is not a power of two >= 2", n);
	if (batch == 0) return SGZ_OK;
	if (uint64_t(n) * batch > SGZ_NOALLOC_MAX_POINTS)
		return fail(SGZ_EINVAL, "sgz_fft_forward_noalloc: n * batch = %llu exceeds the built-in scratch of %u complex points; "
			"split the batch or use the handle API", (unsigned long long)(uint64_t(n) * batch), unsigned(SGZ_NOALLOC_MAX_POINTS));
	if (batch > 65535) return fail(SGZ_EINVAL, "sgz_fft_forward_noalloc: batch %u exceeds 65535", batch);
	{
		// the first pass may write d_out while other threads still read d_in: the two ranges must not overlap at all
		const uintptr_t a = reinterpret_cast<uintptr_t>(d_in), b = reinterpret_cast<uintptr_t>(d_out);
		const uintptr_t bytes = uintptr_t(n) * batch * sizeof(float2);
		if (a < b + bytes && b < a + bytes) return fail(SGZ_EINVAL, "sgz_fft_forward_noalloc: d_in and d_out overlap");
	}

	int dev = -1;
	if (cudaGetDevice(&dev) != cudaSuccess) return fail(SGZ_ECUDA, "sgz_fft_forward_noalloc: no current CUDA device");
	if (dev < 0 || dev >= 64) return fail(SGZ_EUNSUPPORTED, "sgz_fft_forward_noalloc: device ordinal %d out of range", dev);
	if (!g_scratchAddress[dev].load(std::memory_order_acquire))
	{
		// set-up, once per device, with that device current: translate the symbol to its address
		float2* p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noallocScratch);
		if (e != cudaSuccess || !p) return fail(SGZ_ECUDA, "sgz_fft_forward_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
		g_scratchAddress[dev].store(p, std::memory_order_release);
	}
	float2* scratch = g_scratchAddress[dev].load(std::memory_order_acquire);
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);

	uint32_t passes = 0;
	for (uint32_t m = n; m > 1; m >>= 1) ++passes;
	const dim3 block(256), grid((n / 2 + 255) / 256, batch);
	const float2* src = reinterpret_cast<const float2*>(d_in);
	uint32_t len = n, s = 1;
	for (uint32_t i = 1; i <= passes; ++i, len >>= 1, s <<= 1)
	{
		// alternate so that the last pass lands in d_out; every scratch element a pass reads was
		// written by the pass before it, so nothing needs clearing between calls
		float2* dst = ((passes - i) & 1) ? scratch : reinterpret_cast<float2*>(d_out);
		k_fft_pass_noalloc<<<grid, block, 0, stream>>>(src, dst, n, len, s);
		src = dst;
	}
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return fail(SGZ_ECUDA, "sgz_fft_forward_noalloc: launch failed: %s", cudaGetErrorString(e));
	return SGZ_OK;
}
