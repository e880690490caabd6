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
		//     state and leave their final state in a slot (seam[c] if the part began mid-stream, seam[G + c] if it began
		//     the stream); k_seam_carry / k_seam_apply carry the peak hold across afterwards, bit-identically (and
		//     k_phase_smooth runs Phase's second state, which is a filter).  (Round 2 also had the CTAs do this themselves at
		//     the end of the launch: measured no faster than the kernels at any shape, and removed.)
		//     inlineCarry = 1 (a stream in a few parts): k_seam_apply walks the few parts before its own itself, no
		//     k_seam_carry; a part that is a whole stream then stores its final state straight to `states`.
		uint32_t chained, epoch, inlineCarry;
		float2* states;  // [S][2][P]
		float2* seam;    // [2 G][2][P]
		uint32_t* flags; // [G]
		float2* results; // outMode 0: [S][F][2][P]; 1 (LineMain of every frame): [S][F][P]; 2 (last frame only): [S][2][P]
		float2* lastResults; // [S][2][P] lineGraphs[k].results of the handle (see resultRow)
		float2* spare;       // [P] where results that the output mode does not keep are written
		uint32_t outMode;
	};

	// Where graph k of frame f of stream s is written.  Every frame writes both graphs somewhere, so that the store in the
	// display loop is unconditional: what the output mode does not keep goes to `spare` (one row of P points that nobody
	// reads), except that the LAST frame's unkept graphs go straight to the handle's lineGraphs results (`last`, [S][2][P]).
	// The host copies the kept rows of the last frame into `last` afterwards, so `last` is complete in every mode.
	__device__ __forceinline__ float2* resultRow(const FusedArgs& a, uint64_t s, uint32_t f, int k, uint32_t P)
	{
		const uint32_t F = a.nFrames;
		if (a.outMode == SGZ_OUT_FULL) return a.results + ((s * F + f) * 2 + k) * P;
		if (a.outMode == SGZ_OUT_LINE_MAIN) return k == 0 ? a.results + (s * F + f) * P : (f + 1 == F ? a.lastResults + (s * 2 + 1) * P : a.spare);
		return f + 1 == F ? a.results + (s * 2 + k) * P : a.spare;
	}
	// the same for code that revisits results after the kernel (k_seam_apply): nullptr where nothing was kept
	__device__ __forceinline__ float2* keptRow(const FusedArgs& a, uint64_t s, uint32_t f, int k, uint32_t P)
	{
		float2* r = resultRow(a, s, f, k, P);
		return r == a.spare ? nullptr : r;
	}

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

	// mbarrier + bulk copy (SGZ8_BULK_STAGE): one thread asks the copy engine for a contiguous run of this GPU's global
	// memory into this CTA's shared memory; the bytes complete a transaction on the mbarrier the consumers wait on
	__device__ __forceinline__ void mbarInit(void* bar, unsigned count)
	{
		const unsigned b = static_cast<unsigned>(__cvta_generic_to_shared(bar));
		asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(b), "r"(count) : "memory");
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__device__ __forceinline__ void bulkStage(void* smemDst, const void* gmemSrc, unsigned bytes, void* bar)
	{
		const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smemDst)), b = static_cast<unsigned>(__cvta_generic_to_shared(bar));
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(b), "r"(bytes) : "memory");
		asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(d), "l"(gmemSrc), "r"(bytes), "r"(b) : "memory");
	}
	__device__ __forceinline__ void mbarWaitParity(void* bar, unsigned parity)
	{
		const unsigned b = static_cast<unsigned>(__cvta_generic_to_shared(bar));
		unsigned done = 0, spins = 0;
		do
		{
			asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b), "r"(parity) : "memory");
		} while (!done && ++spins < (1u << 22));
		if (!done) __trap(); // a copy that never lands must not hang the GPU
	}

	// seam flags of the chained launch mode (gpu-scope release / acquire)
	__device__ __forceinline__ uint32_t ldAcquire(const uint32_t* p)
	{
		uint32_t v;
		asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
		return v;
	}
	__device__ __forceinline__ void stRelease(uint32_t* p, uint32_t v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }

	// One part of a CTA's range: frames [fBegin, fEnd) of one stream, where its peak-hold state comes from (nullptr: zero)
	// and goes to.  Lives in shared memory, written by thread 0 (planFusedPart) once per part.
	struct FusedPart
	{
		const float2* src; float2* dst;
		const float* audio;   // the stream's first sample
		uint32_t nParts, fBegin, fEnd, publish;
		int stream;
		uint32_t vec4;        // the frame loop may fetch the stream's audio as 16-byte words
		float2* row[2];       // where the current frame's two graphs go (resultRow), renewed by thread 0 every frame
	};
	static_assert(sizeof(FusedPart) == 64, "the fused kernel's shared-memory layout reserves 64 bytes for the part plan");

	// Part q of CTA blockIdx.x (FusedArgs explains the two launch modes).  Also waits for the neighbour's flag when the
	// part continues a stream the previous CTA began.  One thread runs it once per part; everything it computes is dead afterwards.
	__device__ __forceinline__ void planFusedPart(const FusedArgs& a, uint32_t q, uint32_t P, bool wholeWindow, FusedPart* out, uint32_t cta, uint32_t G)
	{
		const uint32_t F = a.nFrames;
		const uint64_t items = uint64_t(a.numStreams) * F;
		const uint64_t i0 = items * cta / G, i1 = items * (cta + 1) / G;
		if (i1 == i0) { out->nParts = 0; return; } // (the host never launches more CTAs than items)
		const uint32_t sA = uint32_t(i0 / F), fA = uint32_t(i0 - uint64_t(sA) * F);
		const uint32_t sZ = uint32_t((i1 - 1) / F), fZ = uint32_t(i1 - uint64_t(sZ) * F);
		const uint32_t nParts = sZ - sA + 1;
		out->nParts = nParts;
		if (q >= nParts) return;
		const uint32_t headFirst = (a.chained && fZ < F && sZ > sA) ? 1u : 0u; // the head of stream sZ runs first and is published
		const uint32_t tailLast = (a.chained && fA != 0) ? 1u : 0u;            // the tail of stream sA runs last, from the published state
		uint32_t sLocal = sA + q;
		if (a.chained)
		{
			if (headFirst && q == 0) sLocal = sZ;
			else
			{
				const uint32_t r = q - headFirst, nMid = nParts - headFirst - tailLast;
				sLocal = r < nMid ? sA + tailLast + r : sA;
			}
		}
		const uint32_t fBegin = sLocal == sA ? fA : 0u, fEnd = sLocal == sZ ? fZ : F;
		const int stream = int(sLocal + a.firstStream);
		out->fBegin = fBegin; out->fEnd = fEnd; out->stream = stream; out->publish = 0;
		const float* sbase = a.audio + uint64_t(stream) * a.streamStride;
		out->audio = sbase;
		out->vec4 = (a.layout == SGZ_LAYOUT_INTERLEAVED && ((reinterpret_cast<uintptr_t>(sbase) & 15) == 0) && ((a.hop & 1) == 0) && wholeWindow) ? 1u : 0u;
		out->src = nullptr;
		if (fBegin == 0) out->src = a.states + uint64_t(stream) * 2 * P;
		else if (a.chained)
		{
			// (published at the very start of the neighbour's work, so this hardly ever spins; the bound turns a lost
			// neighbour - a launch that was not co-resident - into a launch error instead of a hang)
			uint32_t spins = 0;
			while (ldAcquire(a.flags + (cta - 1)) != a.epoch) { __nanosleep(128); if (++spins > (1u << 24)) __trap(); }
			out->src = a.seam + uint64_t(cta - 1) * 2 * P;
		}
		if (a.chained)
		{
			if (fEnd == F) out->dst = a.states + uint64_t(stream) * 2 * P;
			else { out->dst = a.seam + uint64_t(cta) * 2 * P; out->publish = 1; }
		}
		else if (a.inlineCarry && fBegin == 0 && fEnd == F) out->dst = a.states + uint64_t(stream) * 2 * P;
		else out->dst = a.seam + uint64_t(fBegin ? cta : G + cta) * 2 * P;
	}

	// packed bin map entry: bit 31 = max-of-bins, bits 30..16 = count, bits 15..0 = first
	__host__ __device__ __forceinline__ uint32_t packBinMap(int mode, int first, int count) { return (uint32_t(mode) << 31) | (uint32_t(count) << 16) | uint32_t(first); }
