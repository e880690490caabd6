	// ------------------------------------------------------------------------------------------
	// Shared pieces of the fused frame kernel (sgz_fused8.cuh): launch arguments, the display scale, small helpers.
	// (included inside namespace sgz from sgz_spectrum.cu)
	// ------------------------------------------------------------------------------------------
	struct FusedArgs
	{
		const float* audio; uint64_t nSamples; uint64_t streamStride; int layout;
		uint32_t hop, nFrames;
		uint32_t firstStream, numStreams; // the launch covers streams [firstStream, firstStream + numStreams)
		// The launch is a fixed number G of CTAs (at most what the GPU holds at once); CTA c takes the items
		// [c I / G, (c + 1) I / G) of the stream-major list of I = numStreams x nFrames frames, so every CTA does the same
		// amount of work whatever the stream count.  A stream cut by a CTA boundary (a seam) carries its peak-hold state
		// across in one of two ways:
		//   chained = 1 (numStreams >= G, every range holds a whole stream's worth of frames): the CTA that owns the head of
		//     the stream runs it FIRST and publishes the state in seam[c] + flags[c] = epoch; the next CTA runs the tail
		//     LAST, from that state.  Same arithmetic in the same order as one CTA per stream.
		//   chained = 0 (fewer streams than CTAs: a long recording): parts that do not begin a stream start from a zero
		//     state; every part leaves its final state in a slot (seam[c] if it began mid-stream, seam[G + c] if it began
		//     the stream) and k_seam_carry / k_seam_apply carry the hold across afterwards (bit-identical, see there).
		uint32_t chained, epoch;
		float2* states;  // [S][2][P]
		float2* seam;    // [2 G][2][P]
		uint32_t* flags; // [G]
		float2* results; // [S][F][2][P]
	};


	// display scale with the hardware log2 (MUFU.LG2): |error| of lg2.approx is ~2^-22 absolute,
	// i.e. < 1e-7 on the [0, 1] display scale after the multiply by deltaYRecip * ln 2
	__device__ __forceinline__ float fastLogScale(float slope, float st, float minFracRecip, float dyLn2, float clip)
	{
		const float d = slope * st * minFracRecip;
		return d > 0.f ? __log2f(d) * dyLn2 : clip;
	}

	// W_32^k, k = 0..15 (cos, -sin of 2 pi k / 32): the k3-dependent part of the last radix-2 twiddle
	// W_N^(u + T g + 256 k3) = W_N^(u + T g) * W_(N/256)^k3
	__device__ __forceinline__ float2 w32(int k)
	{
		constexpr float c[16] = { 1.f, 0.98078528040323044913f, 0.92387953251128675613f, 0.83146961230254523708f, 0.70710678118654752440f,
			0.55557023301960222474f, 0.38268343236508977173f, 0.19509032201612826785f, 0.f, -0.19509032201612826785f, -0.38268343236508977173f,
			-0.55557023301960222474f, -0.70710678118654752440f, -0.83146961230254523708f, -0.92387953251128675613f, -0.98078528040323044913f };
		constexpr float s[16] = { 0.f, 0.19509032201612826785f, 0.38268343236508977173f, 0.55557023301960222474f, 0.70710678118654752440f,
			0.83146961230254523708f, 0.92387953251128675613f, 0.98078528040323044913f, 1.f, 0.98078528040323044913f, 0.92387953251128675613f,
			0.83146961230254523708f, 0.70710678118654752440f, 0.55557023301960222474f, 0.38268343236508977173f, 0.19509032201612826785f };
		return make_float2(c[k], -s[k]);
	}

	__device__ __forceinline__ float csfTap(const float* csfmag, int i, int N) { return (i >= 0 && i <= N) ? csfmag[i] : 0.f; }

	__device__ __forceinline__ float approxSqrt(float x)
	{
		float r;
		asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); // MUFU.SQRT, <= 1 ulp-class error: far inside the 1e-4 budget
		return r;
	}

	__device__ __forceinline__ void cpAsync16(void* smemDst, const void* gmemSrc)
	{
		const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smemDst));
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gmemSrc) : "memory");
	}
	__device__ __forceinline__ void cpAsyncCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
	__device__ __forceinline__ void cpAsyncWaitAll() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

	// seam flags of the chained launch mode (gpu-scope release / acquire)
	__device__ __forceinline__ uint32_t ldAcquire(const uint32_t* p)
	{
		uint32_t v;
		asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
		return v;
	}
	__device__ __forceinline__ void stRelease(uint32_t* p, uint32_t v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }

	// packed bin map entry: bit 31 = max-of-bins, bits 30..16 = count, bits 15..0 = first
	__host__ __device__ __forceinline__ uint32_t packBinMap(int mode, int first, int count) { return (uint32_t(mode) << 31) | (uint32_t(count) << 16) | uint32_t(first); }
