	// ------------------------------------------------------------------------------------------
	// Resonator bank on the tensor cores (tcgen05 / TMEM) — included inside namespace sgz from sgz_spectrum.cu.
	//
	// The recurrence s <- p s + x[n] (sgz_resonator.cuh; TransformDSP.inl:1215-1295) advanced over a slice of L samples
	// from a zero state is a dot product with the pole's powers,
	//     s_L = sum_j p^(L-1-j) x[j],
	// so the end states of N slices for M filters are ONE real GEMM per component: [M x L] (pole powers, constant per
	// configuration) times [L x N] (the samples, which are real and already "K-major": slice n is L consecutive
	// samples).  That is 4 flop per sample and filter instead of the recurrence's 8, and it runs on tcgen05.mma
	// (kind::tf32) instead of the FP32 pipe.  TF32 keeps 11 significant bits, so both operands are split into a TF32
	// head and a TF32 tail and three products are accumulated (head x head + head x tail + tail x head; the dropped
	// tail x tail term is 2^-22 relative): the accumulation itself is FP32 in TMEM.
	//
	// One CTA (128 threads = the 128 TMEM lanes) owns 128 filter rows and walks tiles of 128 slices x 64 samples
	// (8192 samples of one channel of one stream):
	//   fill   the tile's samples go to shared memory as the B operand (channel matrix, head / tail split on the way),
	//   mma    one thread issues 2 (re, im) x 3 (split terms) x 8 (K steps of 8) tcgen05.mma 128x128x8 into two
	//          accumulators of 128 TMEM columns, then tcgen05.commit -> mbarrier,
	//   chain  thread r reads row r of both accumulators (tcgen05.ld 32x32b) and folds the 128 slice states in time
	//          order, s <- p^L s + d_n, leaving the state after 8192 samples from zero: one `partial` entry per row.
	// k_resonate_combine (sgz_resonator.cuh) then chains the tiles and the stream's real state exactly as it chains the
	// time slices of the FP32 kernels.  The pole-power operand (A) of the CTA's 128 rows stays in shared memory across
	// its tiles; it is built on the host in double precision, already in the shared-memory order of the UMMA
	// "K-major, no swizzle" canonical layout: 8-row x 16-byte core matrices, [k / 4][row / 8][row % 8][k % 4].
	// ------------------------------------------------------------------------------------------
	constexpr int kTcL = 64;                          // samples per slice = GEMM K
	constexpr int kTcSlices = 128;                    // slices per tile   = GEMM N
	constexpr int kTcRows = 128;                      // filter rows per CTA = GEMM M = TMEM lanes
	constexpr int kTcSpan = kTcL * kTcSlices;         // samples per tile (8192)
	constexpr int kTcMatFloats = kTcRows * kTcL;      // one operand matrix (32 KB)
	constexpr uint32_t kTcLbo = (kTcRows / 8) * 128;  // bytes between core matrices along K
	constexpr uint32_t kTcSbo = 128;                  // bytes between core matrices along M / N
	constexpr size_t kTcSmemBytes = size_t(6) * kTcMatFloats * sizeof(float) + 64; // A: re/im x head/tail, B: head/tail, + barrier, TMEM address

	struct ResTcArgs
	{
		const float* audio; uint64_t nSamples; uint64_t streamStride; int layout;
		uint64_t first;      // first sample of the block inside the buffer
		uint32_t count;      // samples in the block
		uint32_t tiles;      // ceil(count / kTcSpan) = time slices per stream in `partial`; the LAST tile ends at the block's end and
		                     // reads zeros where it reaches back into the previous tile's samples (leading zeros leave its sum unchanged)
		uint32_t nStreams;
		const float* aTiles; // [rowTiles][4][kTcMatFloats]: re head, re tail, im head, im tail
		const float2* powL;  // [rowTiles * 128] p^L
	};

	__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

	// UMMA shared-memory descriptor: K-major, no swizzle, version 1 (sm_100)
	__device__ __forceinline__ uint64_t tcDesc(uint32_t addr)
	{
		return uint64_t((addr >> 4) & 0x3FFFu) | (uint64_t(kTcLbo >> 4) << 16) | (uint64_t(kTcSbo >> 4) << 32) | (uint64_t(1) << 46);
	}

	__device__ __forceinline__ float tf32Round(float x)
	{
		uint32_t r;
		asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
		return __uint_as_float(r);
	}

	__global__ void __launch_bounds__(128, 1) k_resonate_tc(DeviceTables t, ResTcArgs a, float2* __restrict__ partial)
	{
		extern __shared__ __align__(128) unsigned char tcSmem[];
		float* aMat = reinterpret_cast<float*>(tcSmem);                    // [4][kTcMatFloats]
		float* bHead = aMat + 4 * kTcMatFloats;                            // [kTcMatFloats]
		float* bTail = bHead + kTcMatFloats;
		uint64_t* bar = reinterpret_cast<uint64_t*>(bTail + kTcMatFloats);
		uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bar + 1);

		const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
		const uint32_t P = t.P, V = t.resVectors, ch = t.resChannels;
		const uint32_t half = P * V, total = half * ch;
		const uint32_t rowTiles = (half + kTcRows - 1) / kTcRows;

		if (warp == 0)
		{
			asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smemAddr(tmemSlot)) : "memory");
			asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
		}
		if (tid == 0)
		{
			asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smemAddr(bar)) : "memory");
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		}
		asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
		__syncthreads();
		asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
		const uint32_t tmem = *tmemSlot;

		// instruction descriptor: D = F32, A = B = TF32, both K-major, N = 128, M = 128
		constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kTcSlices >> 3) << 17) | (uint32_t(kTcRows >> 4) << 24);

		// this CTA's items: (row tile, stream, channel, time tile), row tile slowest so that A is reloaded rarely
		const uint64_t items = uint64_t(rowTiles) * a.nStreams * ch * a.tiles;
		const uint64_t i0 = items * blockIdx.x / gridDim.x, i1 = items * (blockIdx.x + 1) / gridDim.x;
		uint32_t loadedRowTile = 0xFFFFFFFFu, phase = 0;
		for (uint64_t it = i0; it < i1; ++it)
		{
			const uint32_t tt = uint32_t(it % a.tiles);
			const uint32_t c = uint32_t((it / a.tiles) % ch);
			const uint32_t s = uint32_t((it / (uint64_t(a.tiles) * ch)) % a.nStreams);
			const uint32_t rt = uint32_t(it / (uint64_t(a.tiles) * ch * a.nStreams));
			if (rt != loadedRowTile)
			{
				const float4* src = reinterpret_cast<const float4*>(a.aTiles + size_t(rt) * 4 * kTcMatFloats);
				float4* dst = reinterpret_cast<float4*>(aMat);
				for (int i = tid; i < 4 * kTcMatFloats / 4; i += 128) dst[i] = __ldg(src + i);
				loadedRowTile = rt;
			}
			// ---- fill B: slice n = 8 rg + r (row of the operand), samples k = 4 kg + e; a warp task is one rg x four kg.  All
			// of a thread's 64 samples are requested before the first is used: one memory latency per tile, not sixteen ----
			{
				const float* base = a.audio + uint64_t(s) * a.streamStride;
				// samples [lo, hi) of the block belong to this tile; its 8192 slots end at hi
				const uint64_t lo = uint64_t(tt) * kTcSpan, hi = min(lo + uint64_t(kTcSpan), uint64_t(a.count));
				const long long tile0 = (long long)hi - kTcSpan; // block-relative; below lo (even negative) for the block's last, short tile
				const bool whole = tile0 >= (long long)lo;
				const uint32_t r = lane & 7, kq = lane >> 3;
				float2 raw[16][4];
#pragma unroll
				for (int i = 0; i < 16; ++i)
				{
					const uint32_t task = warp + 4u * i, rg = task & 15, kg = (task >> 4) * 4 + kq;
					const long long n0 = tile0 + (long long)((rg * 8 + r) * kTcL + kg * 4);
#pragma unroll
					for (int e = 0; e < 4; ++e)
					{
						raw[i][e] = make_float2(0.f, 0.f);
						if (whole || n0 + e >= (long long)lo) fetchLR(base, a.layout, a.nSamples, a.first + uint64_t(n0 + e), raw[i][e].x, raw[i][e].y);
					}
				}
#pragma unroll
				for (int i = 0; i < 16; ++i)
				{
					const uint32_t task = warp + 4u * i, rg = task & 15, kg = (task >> 4) * 4 + kq;
					float x[4];
#pragma unroll
					for (int e = 0; e < 4; ++e)
					{
						const float2 m = resonatorMatrix(t.conf, raw[i][e].x, raw[i][e].y);
						x[e] = c ? m.y : m.x;
					}
					float4 hd, tl;
					hd.x = tf32Round(x[0]); hd.y = tf32Round(x[1]); hd.z = tf32Round(x[2]); hd.w = tf32Round(x[3]);
					tl.x = tf32Round(x[0] - hd.x); tl.y = tf32Round(x[1] - hd.y); tl.z = tf32Round(x[2] - hd.z); tl.w = tf32Round(x[3] - hd.w);
					const uint32_t off = ((kg * 16 + rg) * 8 + r) * 4;
					*reinterpret_cast<float4*>(bHead + off) = hd;
					*reinterpret_cast<float4*>(bTail + off) = tl;
				}
			}
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the operands were written through the generic proxy
			__syncthreads();
			// ---- mma: D_re = A_re X, D_im = A_im X, each as head x head + head x tail + tail x head ----
			if (tid == 0)
			{
				asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
				const uint32_t aBase = smemAddr(aMat), bH = smemAddr(bHead), bT = smemAddr(bTail);
#pragma unroll
				for (int comp = 0; comp < 2; ++comp)
				{
					const uint32_t d = tmem + uint32_t(comp) * kTcSlices;
					const uint32_t aH = aBase + uint32_t(2 * comp) * kTcMatFloats * 4, aT = aH + kTcMatFloats * 4;
#pragma unroll
					for (int term = 0; term < 3; ++term)
					{
						const uint32_t aAddr = term == 2 ? aT : aH, bAddr = term == 1 ? bT : bH;
#pragma unroll
						for (int ks = 0; ks < kTcL / 8; ++ks)
						{
							const uint64_t da = tcDesc(aAddr + uint32_t(ks) * 2 * kTcLbo), db = tcDesc(bAddr + uint32_t(ks) * 2 * kTcLbo);
							const uint32_t accumulate = (term | ks) ? 1u : 0u;
							asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
								:: "r"(d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
						}
					}
				}
				asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smemAddr(bar)) : "memory");
			}
			// ---- wait for the accumulators ----
			{
				uint32_t done = 0;
				while (!done)
				{
					asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
						: "=r"(done) : "r"(smemAddr(bar)), "r"(phase) : "memory");
				}
				phase ^= 1u;
			}
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
			// ---- chain the 128 slice states of row g in time order ----
			{
				const uint32_t g = rt * kTcRows + tid;
				const float2 pL = a.powL[g];
				float2 st = make_float2(0.f, 0.f);
				const uint32_t laneBase = tmem + ((warp * 32u) << 16);
#pragma unroll 1
				for (int cb = 0; cb < kTcSlices / 16; ++cb)
				{
					uint32_t re[16], im[16];
					asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
						: "=r"(re[0]), "=r"(re[1]), "=r"(re[2]), "=r"(re[3]), "=r"(re[4]), "=r"(re[5]), "=r"(re[6]), "=r"(re[7]),
						  "=r"(re[8]), "=r"(re[9]), "=r"(re[10]), "=r"(re[11]), "=r"(re[12]), "=r"(re[13]), "=r"(re[14]), "=r"(re[15])
						: "r"(laneBase + uint32_t(cb) * 16));
					asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
						: "=r"(im[0]), "=r"(im[1]), "=r"(im[2]), "=r"(im[3]), "=r"(im[4]), "=r"(im[5]), "=r"(im[6]), "=r"(im[7]),
						  "=r"(im[8]), "=r"(im[9]), "=r"(im[10]), "=r"(im[11]), "=r"(im[12]), "=r"(im[13]), "=r"(im[14]), "=r"(im[15])
						: "r"(laneBase + uint32_t(kTcSlices) + uint32_t(cb) * 16));
					asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
					for (int j = 0; j < 16; ++j)
					{
						const float nr = st.x * pL.x - st.y * pL.y + __uint_as_float(re[j]);
						const float ni = st.x * pL.y + st.y * pL.x + __uint_as_float(im[j]);
						st = make_float2(nr, ni);
					}
				}
				if (g < half) partial[(uint64_t(s) * a.tiles + tt) * total + uint64_t(c) * half + g] = st;
			}
			asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
			__syncthreads(); // the accumulators and the B operand are free again
		}
		__syncthreads();
		if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tmem) : "memory");
	}
