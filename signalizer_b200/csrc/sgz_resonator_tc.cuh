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
	// One CTA owns 128 filter rows (= the 128 TMEM lanes) and walks tiles of 64 slices x 64 samples (4096 samples of one
	// channel of one stream), its warps in three roles that overlap through mbarriers (two stages in flight):
	//   fill   (8 warps) the tile's samples go to shared memory as the B operand (channel matrix, head / tail split),
	//   mma    (1 thread) 2 (re, im) x 3 (split terms) x 8 (K steps of 8) tcgen05.mma 128x64x8 into two accumulators of
	//          64 TMEM columns, then tcgen05.commit -> mbarriers,
	//   chain  (4 warps) thread r reads row r of both accumulators (tcgen05.ld 32x32b) and folds the 64 slice states in
	//          time order, s <- p^L s + d_n, leaving the state after 4096 samples from zero: one `partial` entry per row.
	// k_resonate_combine (sgz_resonator.cuh) then chains the tiles and the stream's real state exactly as it chains the
	// time slices of the FP32 kernels.  The pole-power operand (A) of the CTA's 128 rows stays in shared memory across
	// its tiles; it is built on the host in double precision, already in the shared-memory order of the UMMA
	// "K-major, no swizzle" canonical layout: 8-row x 16-byte core matrices, [k / 4][row / 8][row % 8][k % 4].
	// ------------------------------------------------------------------------------------------
#ifndef SGZ_TC_A_TMEM
#define SGZ_TC_A_TMEM 0 // 1 = the pole-power operand lives in tensor memory (columns 256..511) instead of shared memory, so that an MMA reads 2 KB of
                        // shared memory, not 6.  Measured (B200): same results, same time (0.120 vs 0.117 ms on BASELINE config 2) - the 125 cycles a
                        // 128x64x8 TF32 MMA takes here are not operand fetch
#endif
	constexpr int kTcL = 64;                          // samples per slice = GEMM K
	constexpr int kTcSlices = 64;                     // slices per tile   = GEMM N
	constexpr int kTcRows = 128;                      // filter rows per CTA = GEMM M = TMEM lanes
	constexpr int kTcSpan = kTcL * kTcSlices;         // samples per tile (4096)
	constexpr int kTcAFloats = kTcRows * kTcL;        // one A matrix (32 KB)
	constexpr int kTcBFloats = kTcSlices * kTcL;      // one B matrix (16 KB)
	constexpr int kTcStages = 2;                      // B operand and accumulator stages in flight
	constexpr uint32_t kTcSbo = 128;                  // bytes between core matrices along M / N
	constexpr uint32_t kTcLboA = (kTcRows / 8) * 128; // bytes between core matrices along K, A operand
	constexpr uint32_t kTcLboB = (kTcSlices / 8) * 128; // ... B operand
	constexpr uint32_t kTcTmemColumns = SGZ_TC_A_TMEM ? 512 : 256; // accumulators: 2 stages x (re, im) x 64; then A: (re, im) x (head, tail) x 64
	constexpr int kTcChainWarps = 4, kTcFillWarps = 8;
	constexpr int kTcThreads = 32 * (kTcChainWarps + kTcFillWarps + 1);
	constexpr size_t kTcSmemBytes = size_t(4) * kTcAFloats * 4 + size_t(kTcStages) * 2 * kTcBFloats * 4 + 128; // A: re/im x head/tail; B: stages x head/tail; barriers, TMEM address

	struct ResTcArgs
	{
		const float* audio; uint64_t nSamples; uint64_t streamStride; int layout;
		uint64_t first;      // first sample of the block inside the buffer
		uint32_t count;      // samples in the block
		uint32_t tiles;      // ceil(count / kTcSpan) = time slices per stream in `partial`; the LAST tile ends at the block's end and
		                     // reads zeros where it reaches back into the previous tile's samples (leading zeros leave its sum unchanged)
		uint32_t nStreams;
		const float* aTiles; // [rowTiles][4][kTcAFloats]: re head, re tail, im head, im tail
		const float2* powL;  // [rowTiles * 128] p^L
	};

	__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

	// UMMA shared-memory descriptor: K-major, no swizzle, version 1 (sm_100)
	__device__ __forceinline__ uint64_t tcDesc(uint32_t addr, uint32_t lbo)
	{
		return uint64_t((addr >> 4) & 0x3FFFu) | (uint64_t(lbo >> 4) << 16) | (uint64_t(kTcSbo >> 4) << 32) | (uint64_t(1) << 46);
	}

	__device__ __forceinline__ float tf32Round(float x)
	{
		uint32_t r;
		asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
		return __uint_as_float(r);
	}

	__device__ __forceinline__ void mbarInit(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smemAddr(bar)), "r"(count) : "memory"); }
	__device__ __forceinline__ void mbarArrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smemAddr(bar)) : "memory"); }
	__device__ __forceinline__ void mbarWait(uint64_t* bar, uint32_t parity)
	{
		uint32_t done = 0;
		while (!done)
			asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
				: "=r"(done) : "r"(smemAddr(bar)), "r"(parity) : "memory");
	}
	__device__ __forceinline__ void tcCommit(uint64_t* bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smemAddr(bar)) : "memory"); }

	// item -> (row tile, stream, channel, time tile); the row tile is the slowest index so that the A operand is reloaded rarely
	struct TcItem { uint32_t rt, s, c, tt; };
	__device__ __forceinline__ TcItem tcItem(uint64_t it, uint32_t tiles, uint32_t ch, uint32_t nStreams)
	{
		TcItem r;
		r.tt = uint32_t(it % tiles);
		r.c = uint32_t((it / tiles) % ch);
		r.s = uint32_t((it / (uint64_t(tiles) * ch)) % nStreams);
		r.rt = uint32_t(it / (uint64_t(tiles) * ch * nStreams));
		return r;
	}
	// the item after q (the 64-bit divisions of tcItem once per segment, not once per tile)
	__device__ __forceinline__ void tcNext(TcItem& q, uint32_t tiles, uint32_t ch, uint32_t nStreams)
	{
		if (++q.tt < tiles) return;
		q.tt = 0;
		if (++q.c < ch) return;
		q.c = 0;
		if (++q.s < nStreams) return;
		q.s = 0; ++q.rt;
	}

	// Warp roles (13 warps): 0-3 chain (they own the four TMEM lane quarters), 4-11 fill, 12 issues the MMAs.  Two stages of
	// the B operand and of the accumulators are in flight, handed over through mbarriers:
	//   bFull[s]  fill -> mma    (8 arrivals, one per fill warp)       bEmpty[s]  mma -> fill   (tcgen05.commit)
	//   tFull[s]  mma  -> chain  (tcgen05.commit)                      tEmpty[s]  chain -> mma  (4 arrivals, one per chain warp)
	// so the tensor cores work on tile i while tile i + 1 is being filled and tile i - 1 chained.
	__global__ void __launch_bounds__(kTcThreads, 1) k_resonate_tc(DeviceTables t, ResTcArgs a, float2* __restrict__ partial)
	{
		extern __shared__ __align__(128) unsigned char tcSmem[];
		float* aMat = reinterpret_cast<float*>(tcSmem);                              // [4][kTcAFloats]
		float* bMat = aMat + 4 * kTcAFloats;                                         // [stage][head / tail][kTcBFloats]
		uint64_t* bars = reinterpret_cast<uint64_t*>(bMat + kTcStages * 2 * kTcBFloats);
		uint64_t* bFull = bars, *bEmpty = bars + kTcStages, *tFull = bars + 2 * kTcStages, *tEmpty = bars + 3 * kTcStages;
		uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bars + 4 * kTcStages);

		const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
		const uint32_t P = t.P, V = t.resVectors, ch = t.resChannels;
		const uint32_t half = P * V, total = half * ch;
		const uint32_t rowTiles = (half + kTcRows - 1) / kTcRows;

		if (warp == 0)
		{
			asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smemAddr(tmemSlot)), "r"(kTcTmemColumns) : "memory");
			asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
		}
		if (tid == 0)
		{
			for (int s = 0; s < kTcStages; ++s) { mbarInit(bFull + s, kTcFillWarps); mbarInit(bEmpty + s, 1); mbarInit(tFull + s, 1); mbarInit(tEmpty + s, kTcChainWarps); }
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		}
		asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
		__syncthreads();
		asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
		const uint32_t tmem = *tmemSlot;

		// instruction descriptor: D = F32, A = B = TF32, both K-major, N = 64, M = 128
		constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kTcSlices >> 3) << 17) | (uint32_t(kTcRows >> 4) << 24);

		const uint64_t items = uint64_t(rowTiles) * a.nStreams * ch * a.tiles;
		const uint64_t i0 = items * blockIdx.x / gridDim.x, i1 = items * (blockIdx.x + 1) / gridDim.x;
		const uint64_t perRowTile = uint64_t(a.tiles) * ch * a.nStreams;
		uint32_t n = 0; // tiles this CTA has started: stage = n & 1, use = n >> 1 (the same count in every role)
		for (uint64_t seg0 = i0; seg0 < i1;)
		{
			// ---- a segment: the CTA's items of one row tile; its A operand is loaded by everybody, then the roles part ----
			const uint32_t rt = uint32_t(seg0 / perRowTile);
			const uint64_t seg1 = min(i1, uint64_t(rt + 1) * perRowTile);
#if SGZ_TC_A_TMEM
			if (warp < kTcChainWarps)
			{
				// thread = TMEM lane = filter row: its 64 pole powers of each of the four matrices go to columns 256 + 64 mi + k.
				// The host tiles are in the shared-memory operand order [k / 4][row / 8][row % 8][k % 4]: for a given k / 4 the
				// rows are consecutive 16-byte words, so a warp's loads are contiguous.
				const float4* src = reinterpret_cast<const float4*>(a.aTiles + size_t(rt) * 4 * kTcAFloats);
				const uint32_t laneBase = tmem + ((warp * 32u) << 16);
				for (int mi = 0; mi < 4; ++mi)
#pragma unroll 1
					for (int kc = 0; kc < 4; ++kc)
					{
						float4 v[4];
#pragma unroll
						for (int j = 0; j < 4; ++j) v[j] = __ldg(src + size_t(mi) * (kTcAFloats / 4) + size_t(kc * 4 + j) * kTcRows + tid);
						asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
							:: "r"(laneBase + 256u + uint32_t(mi) * 64u + uint32_t(kc) * 16u),
							   "r"(__float_as_uint(v[0].x)), "r"(__float_as_uint(v[0].y)), "r"(__float_as_uint(v[0].z)), "r"(__float_as_uint(v[0].w)),
							   "r"(__float_as_uint(v[1].x)), "r"(__float_as_uint(v[1].y)), "r"(__float_as_uint(v[1].z)), "r"(__float_as_uint(v[1].w)),
							   "r"(__float_as_uint(v[2].x)), "r"(__float_as_uint(v[2].y)), "r"(__float_as_uint(v[2].z)), "r"(__float_as_uint(v[2].w)),
							   "r"(__float_as_uint(v[3].x)), "r"(__float_as_uint(v[3].y)), "r"(__float_as_uint(v[3].z)), "r"(__float_as_uint(v[3].w)) : "memory");
					}
				asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
				asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
			}
			__syncthreads();
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#else
			{
				const float4* src = reinterpret_cast<const float4*>(a.aTiles + size_t(rt) * 4 * kTcAFloats);
				float4* dst = reinterpret_cast<float4*>(aMat);
				for (int i = tid; i < 4 * kTcAFloats / 4; i += kTcThreads) dst[i] = __ldg(src + i);
				asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
			}
			__syncthreads();
#endif
			if (warp >= kTcChainWarps && warp < kTcChainWarps + kTcFillWarps)
			{
				// ---- fill: operand row = slice n = 8 rg + r, k = 4 kg + e; a warp task is one rg x four kg; a thread's 16
				// samples are requested before the first is used ----
				const uint32_t fw = warp - kTcChainWarps, r = lane & 7, kq = lane >> 3;
				uint32_t m = n;
				TcItem q = tcItem(seg0, a.tiles, ch, a.nStreams);
				for (uint64_t it = seg0; it < seg1; ++it, ++m, tcNext(q, a.tiles, ch, a.nStreams))
				{
					const uint32_t st = m & 1u;
					const float* base = a.audio + uint64_t(q.s) * a.streamStride;
					const uint64_t lo = uint64_t(q.tt) * kTcSpan, hi = min(lo + uint64_t(kTcSpan), uint64_t(a.count));
					const long long tile0 = (long long)hi - kTcSpan; // block-relative; below lo (even negative) for the block's last, short tile
					const bool whole = tile0 >= (long long)lo;
					float2 raw[4][4];
					// (a whole tile of an interleaved, 16-byte aligned stream: a thread's four consecutive samples are two 16-byte words)
					const bool wide = whole && a.layout == SGZ_LAYOUT_INTERLEAVED && (reinterpret_cast<uintptr_t>(base) & 15) == 0 &&
						((a.first + uint64_t(tile0)) & 1) == 0;
#pragma unroll
					for (int i = 0; i < 4; ++i)
					{
						const uint32_t task = fw + kTcFillWarps * i, rg = task & 7, kg = (task >> 3) * 4 + kq;
						const long long n0 = tile0 + (long long)((rg * 8 + r) * kTcL + kg * 4);
						if (wide)
						{
							const float4* p = reinterpret_cast<const float4*>(base + 2 * (a.first + uint64_t(n0)));
							const float4 v0 = __ldg(p), v1 = __ldg(p + 1);
							raw[i][0] = make_float2(v0.x, v0.y); raw[i][1] = make_float2(v0.z, v0.w);
							raw[i][2] = make_float2(v1.x, v1.y); raw[i][3] = make_float2(v1.z, v1.w);
						}
						else
						{
#pragma unroll
							for (int e = 0; e < 4; ++e)
							{
								raw[i][e] = make_float2(0.f, 0.f);
								if (whole || n0 + e >= (long long)lo) fetchLR(base, a.layout, a.nSamples, a.first + uint64_t(n0 + e), raw[i][e].x, raw[i][e].y);
							}
						}
					}
					mbarWait(bEmpty + st, ((m >> 1) & 1u) ^ 1u); // the MMAs that read this stage two tiles ago are done
					float* bHead = bMat + size_t(st) * 2 * kTcBFloats, *bTail = bHead + kTcBFloats;
#pragma unroll
					for (int i = 0; i < 4; ++i)
					{
						const uint32_t task = fw + kTcFillWarps * i, rg = task & 7, kg = (task >> 3) * 4 + kq;
						float x[4];
#pragma unroll
						for (int e = 0; e < 4; ++e)
						{
							const float2 mx = resonatorMatrix(t.conf, raw[i][e].x, raw[i][e].y);
							x[e] = q.c ? mx.y : mx.x;
						}
						float4 hd, tl;
						hd.x = tf32Round(x[0]); hd.y = tf32Round(x[1]); hd.z = tf32Round(x[2]); hd.w = tf32Round(x[3]);
						tl.x = tf32Round(x[0] - hd.x); tl.y = tf32Round(x[1] - hd.y); tl.z = tf32Round(x[2] - hd.z); tl.w = tf32Round(x[3] - hd.w);
						const uint32_t off = ((kg * (kTcSlices / 8) + rg) * 8 + r) * 4;
						*reinterpret_cast<float4*>(bHead + off) = hd;
						*reinterpret_cast<float4*>(bTail + off) = tl;
					}
					asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // written through the generic proxy, read by the tensor cores
					__syncwarp();
					if (lane == 0) mbarArrive(bFull + st);
				}
			}
			else if (warp == kTcChainWarps + kTcFillWarps)
			{
				// ---- mma: D_re = A_re X, D_im = A_im X, each as head x head + head x tail + tail x head ----
				uint32_t m = n;
				for (uint64_t it = seg0; it < seg1; ++it, ++m)
				{
					const uint32_t st = m & 1u, use = (m >> 1) & 1u;
					mbarWait(bFull + st, use);
					mbarWait(tEmpty + st, use ^ 1u); // the chain has read this accumulator stage
					asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
					if (lane == 0)
					{
						[[maybe_unused]] const uint32_t aBase = smemAddr(aMat);
						const uint32_t bH = smemAddr(bMat + size_t(st) * 2 * kTcBFloats), bT = bH + kTcBFloats * 4;
#pragma unroll
						for (int comp = 0; comp < 2; ++comp)
						{
							const uint32_t d = tmem + st * (2 * kTcSlices) + uint32_t(comp) * kTcSlices;
#if !SGZ_TC_A_TMEM
							const uint32_t aH = aBase + uint32_t(2 * comp) * kTcAFloats * 4, aT = aH + kTcAFloats * 4;
#endif
#pragma unroll
							for (int term = 0; term < 3; ++term)
							{
								const uint32_t bAddr = term == 1 ? bT : bH;
#if !SGZ_TC_A_TMEM
								const uint32_t aAddr = term == 2 ? aT : aH;
#endif
#pragma unroll
								for (int ks = 0; ks < kTcL / 8; ++ks)
								{
									const uint64_t db = tcDesc(bAddr + uint32_t(ks) * 2 * kTcLboB, kTcLboB);
									const uint32_t accumulate = (term | ks) ? 1u : 0u;
#if SGZ_TC_A_TMEM
									// A from tensor memory: matrix 2 comp (+ 1 for the tail), eight columns per K step
									const uint32_t ta = tmem + 256u + uint32_t(2 * comp + (term == 2 ? 1 : 0)) * 64u + uint32_t(ks) * 8u;
									asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
										:: "r"(d), "r"(ta), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
#else
									const uint64_t da = tcDesc(aAddr + uint32_t(ks) * 2 * kTcLboA, kTcLboA);
									asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
										:: "r"(d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
#endif
								}
							}
						}
						tcCommit(bEmpty + st); // the B stage may be refilled ...
						tcCommit(tFull + st);  // ... and the accumulators read, once these MMAs have completed
					}
					__syncwarp();
				}
			}
			else
			{
				// ---- chain (warps 0-3): thread = TMEM lane = filter row; fold the tile's 64 slice states in time order ----
				const uint32_t g = rt * kTcRows + tid;
				const float2 pL = a.powL[g];
				const uint32_t laneBase = tmem + ((warp * 32u) << 16);
				uint32_t m = n;
				TcItem q = tcItem(seg0, a.tiles, ch, a.nStreams);
				for (uint64_t it = seg0; it < seg1; ++it, ++m, tcNext(q, a.tiles, ch, a.nStreams))
				{
					const uint32_t st = m & 1u;
					mbarWait(tFull + st, (m >> 1) & 1u);
					asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
					float2 state = make_float2(0.f, 0.f);
#pragma unroll 1
					for (int cb = 0; cb < kTcSlices / 16; ++cb)
					{
						uint32_t re[16], im[16];
						asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
							: "=r"(re[0]), "=r"(re[1]), "=r"(re[2]), "=r"(re[3]), "=r"(re[4]), "=r"(re[5]), "=r"(re[6]), "=r"(re[7]),
							  "=r"(re[8]), "=r"(re[9]), "=r"(re[10]), "=r"(re[11]), "=r"(re[12]), "=r"(re[13]), "=r"(re[14]), "=r"(re[15])
							: "r"(laneBase + st * (2 * kTcSlices) + uint32_t(cb) * 16));
						asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
							: "=r"(im[0]), "=r"(im[1]), "=r"(im[2]), "=r"(im[3]), "=r"(im[4]), "=r"(im[5]), "=r"(im[6]), "=r"(im[7]),
							  "=r"(im[8]), "=r"(im[9]), "=r"(im[10]), "=r"(im[11]), "=r"(im[12]), "=r"(im[13]), "=r"(im[14]), "=r"(im[15])
							: "r"(laneBase + st * (2 * kTcSlices) + uint32_t(kTcSlices) + uint32_t(cb) * 16));
						asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
						if (cb == kTcSlices / 16 - 1)
						{
							// everything of this stage is in registers: hand the accumulators back before the arithmetic
							asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
							__syncwarp();
							if (lane == 0) mbarArrive(tEmpty + st);
						}
#pragma unroll
						for (int j = 0; j < 16; ++j)
						{
							const float nr = state.x * pL.x - state.y * pL.y + __uint_as_float(re[j]);
							const float ni = state.x * pL.y + state.y * pL.x + __uint_as_float(im[j]);
							state = make_float2(nr, ni);
						}
					}
					if (g < half) partial[(uint64_t(q.s) * a.tiles + q.tt) * total + uint64_t(q.c) * half + g] = state;
				}
			}
			n += uint32_t(seg1 - seg0);
			seg0 = seg1;
			__syncthreads(); // every role is through the segment (the chain has seen the last accumulators): A may be replaced
		}
		asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
		__syncthreads();
		if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(kTcTmemColumns) : "memory");
	}
