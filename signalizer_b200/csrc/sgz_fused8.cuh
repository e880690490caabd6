	// ------------------------------------------------------------------------------------------
	// k_fused8<K1> — the fused frame kernel with 64 K1 threads per stream and 8 (E, O) pairs per thread: K1 = 8 is
	// N = 8192 (described below), K1 = 4 is N = 4096 (same passes 2-4, a radix-4 first pass; see Fused8GeomT) (included inside namespace sgz from sgz_spectrum.cu).
	//
	// Same arithmetic and shared-memory budget as k_fused<16,...>, different shape: M = 4096 = 8^4, four
	// register-resident radix-8 passes, 32 data registers per thread instead of 64, so that the kernel
	// fits 64 registers and TWO 512-thread CTAs (32 warps) are resident per SM.  ncu on k_fused showed
	// no pipe saturated and 16 warps/SM unable to cover shared-memory / L2 latency; this variant buys
	// latency tolerance with warps instead of instruction-level parallelism.
	//
	//   n = n1 512 + n2 64 + n3 8 + n4,  k = k1 + 8 k2 + 64 k3 + 512 k4          (all digits 0..7)
	//   pass 1: thread (n2 n3 n4) DFT8 over n1, twiddle W_M^((64 n2 + v) k1) = W_64^(n2 k1) W_M^(v k1), v = 8 n3 + n4
	//   pass 2: thread (k1 n3 n4) DFT8 over n2, twiddle W_512^(v k2)
	//   pass 3: thread (k1 k2 n4) DFT8 over n3, twiddle W_64^(n4 k3)
	//   pass 4: thread (k3 k2 k1) DFT8 over n4  -> bins k = t + 512 k4, t = k1 + 8 k2 + 64 k3
	// exchange element (k1, d2, d3, d4) lives at k1 512 + d2 64 + d3 8 + (d4 ^ k1): every pass reads and
	// writes 128-bit words without bank conflicts and passes 2-3 run in place.
	// ------------------------------------------------------------------------------------------
// SGZ8_TW_POWERS: 0 = every twiddle from the shared-memory tables, 1 = passes 1-2 derive w^2..w^7 from w^1 in
// registers (measured +2.0 % on BASELINE config 3), 2 = pass 3 as well (no further gain; tools/ab_bench.sh)
#ifndef SGZ8_TW_POWERS
#define SGZ8_TW_POWERS 1
#endif
// SGZ8_WINDOW: 0 = window taps from the table, fetched ahead of the audio; 1 = fetched after it; 2 = evaluated from
// the thread's twiddle when the window is a cosine sum spanning the whole transform (table otherwise)
#ifndef SGZ8_WINDOW
#define SGZ8_WINDOW 1
#endif
// A/B switches (tools/ab_libs.sh): SGZ8_SPLIT_SMEM 1 = the two-for-one split fetches its partner bins through shared
// memory instead of warp shuffles; SGZ8_STAGE_LATE 1 = the next frame's audio is staged after the split instead of
// before it; SGZ8_AOS_WEIGHTS 1 = Lanczos weights read point-major
#ifndef SGZ8_SPLIT_SMEM
#define SGZ8_SPLIT_SMEM 1
#endif
// SGZ8_PAIR_BARRIER 1 = the barrier between passes 2 and 3 covers only the two warps that share a sub-transform
#ifndef SGZ8_TW_BROADCAST
#define SGZ8_TW_BROADCAST 0 // twiddle powers derived as broadcast pairs with packed arithmetic (no scalar products, no pairing moves)
#endif
// SGZ8_PHASE_CLOCK (diagnostic build only): thread 0 of CTA 0 adds up the SM clock between the phase boundaries of its
// frames and prints the sums when the kernel ends - where a CTA's frame time goes (tools/ab_libs.sh build pc "-DSGZ8_PHASE_CLOCK")
#ifdef SGZ8_PHASE_CLOCK
#define SGZ8_TICK(i) phaseTick(i)
#else
#define SGZ8_TICK(i)
#endif
#ifndef SGZ8_DUAL
#define SGZ8_DUAL 0 // the headline shape runs two ranges per 1024-thread CTA (k_fused8<..., DUAL = true>)
#endif
#ifndef SGZ8_DUAL_TURNS
#define SGZ8_DUAL_TURNS 0 // ... and the two halves take turns at the exchange-heavy part of a frame
#endif
#ifndef SGZ8_BULK_STAGE
#define SGZ8_BULK_STAGE 1 // next frame's newest half of the audio: 1 = one cp.async.bulk by thread 0 onto an mbarrier instead of four cp.async per thread
#endif
#ifndef SGZ8_SCAN_ROTATE
#define SGZ8_SCAN_ROTATE 0 // max-of-bins scan: lanes start their runs at different offsets so that even run lengths do not meet in one bank
#endif
#ifndef SGZ8_PAIR_BARRIER
#define SGZ8_PAIR_BARRIER 0
#endif
#ifndef SGZ8_STAGE_LATE
#define SGZ8_STAGE_LATE 0
#endif
#ifndef SGZ8_AOS_WEIGHTS
#define SGZ8_AOS_WEIGHTS 0
#endif
	// K1 = radix of the first pass: 8 -> N = 8192 (M = 8^4, 512 threads); 4 -> N = 4096 (M = 4 x 8^3, 256 threads, whose
	// first pass is two radix-4 butterflies per thread); 2 -> N = 2048 (128 threads, four radix-2 butterflies); 1 -> N = 1024
	// (64 threads, no first pass: M = 512 is the sub-transform itself); 16 -> N = 16384 (1024 threads, the radix-16 first
	// pass split over thread pairs with one extra exchange); passes 2-4 and everything after are the same code
	template<int K1>
	struct Fused8GeomT
	{
		static constexpr int T = 64 * K1, M = 512 * K1, N = 2 * M;
		static constexpr size_t xbufBytes = size_t(M) * 16;
		static constexpr size_t twOffset = xbufBytes + 16;
#if SGZ8_TW_POWERS >= 1
		// with the powers derived in registers only A[8][8] (pass 3) and row 1 of C (the base W_512^v of pass 2) are
		// read: 1 KB instead of 8.5 KB per CTA, which goes back to the L1 that the window and weight tables live in
		static constexpr int twEntries = 64 + 64;
#else
		static constexpr int twEntries = 64 + 512 + 512;            // A[8][8] | B[8][64] | C[8][64]
#endif
		static constexpr size_t twBytes = size_t(twEntries) * sizeof(float2);
		static constexpr size_t stateOffset = twOffset + twBytes;
		__host__ __device__ static size_t mapOffset(uint32_t P) { return stateOffset + size_t(P) * 16; }
		__host__ __device__ static size_t partOffset(uint32_t P) { return (mapOffset(P) + size_t(P) * 12 + 15) & ~size_t(15); }
		__host__ __device__ static size_t smemBytes(uint32_t P) { return partOffset(P) + 80; } // ... + the part plan (FusedPart, 64 bytes) + the staging mbarrier
	};
	typedef Fused8GeomT<8> Fused8Geom;

	// 8-point forward DFT on SoA pairs; X[k] ends up in v[perm8(k)]
	__device__ __forceinline__ void dft8q(float2 (&re)[8], float2 (&im)[8])
	{
		constexpr float R = 0.70710678118654752440f;
		radix4p(re[0], im[0], re[2], im[2], re[4], im[4], re[6], im[6]);
		radix4p(re[1], im[1], re[3], im[3], re[5], im[5], re[7], im[7]);
		{
			const float2 a = add2(re[3], im[3]), b = sub2(im[3], re[3]);   // W8^1
			re[3] = mul2(a, bc2(R)); im[3] = mul2(b, bc2(R));
			const float2 t = re[5]; re[5] = im[5]; im[5] = sub2(bc2(0.f), t); // W8^2 = -i
			const float2 c = sub2(im[7], re[7]), d = add2(re[7], im[7]);   // W8^3
			re[7] = mul2(c, bc2(R)); im[7] = mul2(d, bc2(-R));
		}
#pragma unroll
		for (int m = 0; m < 4; ++m)
		{
			const float2 ar = re[2 * m], ai = im[2 * m], br = re[2 * m + 1], bi = im[2 * m + 1];
			re[2 * m] = add2(ar, br); im[2 * m] = add2(ai, bi);
			re[2 * m + 1] = sub2(ar, br); im[2 * m + 1] = sub2(ai, bi);
		}
	}

#ifdef SGZ8_CTA_CLOCK
	__device__ __forceinline__ unsigned long long& ctaClock0() { __shared__ unsigned long long b; return b; }
#endif
#ifdef SGZ8_PHASE_CLOCK
	__device__ __forceinline__ long long* phaseClockBuffer() { __shared__ long long b[18]; return b; } // [0..15] sums, [16] last clock, [17] frames
	__device__ __forceinline__ void phaseTick(int i)
	{
		if (threadIdx.x != 0 || blockIdx.x != 0) return;
		long long* b = phaseClockBuffer();
		const long long c = clock64();
		if (i >= 0) b[i] += c - b[16];
		b[16] = c;
	}
#endif
	// exchange element (k1, d2, d3, d4) of the 512-point sub-transform k1.  The XOR keeps every pass free of bank
	// conflicts on 128-bit words: the lanes of a quarter warp differ in d4 (passes 1-3) or, in pass 4, in k1 and - when
	// k1 has fewer than three bits (K1 = 4, 2) - in the low bit(s) of d2 as well
	template<int K1>
	__device__ __forceinline__ int phys8(int k1, int d2, int d3, int d4)
	{
		return k1 * 512 + d2 * 64 + d3 * 8 + (d4 ^ (K1 >= 8 ? (k1 & 7) : K1 == 4 ? (k1 | ((d2 & 1) << 2)) : K1 == 2 ? (k1 | ((d2 & 3) << 1)) : (d2 & 7)));
	}

	// Multiplies output k of a radix-8 butterfly (held in slot perm8(k)) by w1^k, k = 1..7.  The powers are
	// derived in registers (depth <= 3 complex products: w2 = w1^2, w3 = w2 w1, w4 = w2^2, w5 = w4 w1,
	// w6 = w3^2, w7 = w4 w3) instead of being read from a [8][64] table: the kernel is bound by the
	// shared-memory pipe, not by FP issue, and this trades 7 LDS.64 for 6 complex products.
	__device__ __forceinline__ void applyTwiddlePowers(float2 w1, float2 (&re)[8], float2 (&im)[8])
	{
#if SGZ8_TW_BROADCAST
		// the powers are derived as (x, x), (y, y) pairs, the form the packed multiply wants: five packed operations per
		// product instead of four scalar ones and the moves that pair their results up
		const float2 X1 = bc2(w1.x), Y1 = bc2(w1.y);
		auto product = [](float2 Xa, float2 Ya, float2 Xb, float2 Yb, float2& X, float2& Y)
		{
			X = sub2(mul2(Xa, Xb), mul2(Ya, Yb));
			Y = fma2(Xa, Yb, mul2(Ya, Xb));
		};
		auto apply = [&](int k, float2 X, float2 Y)
		{
			float2& r = re[perm8(k)]; float2& i = im[perm8(k)];
			const float2 nr = sub2(mul2(r, X), mul2(i, Y));
			const float2 ni = fma2(r, Y, mul2(i, X));
			r = nr; i = ni;
		};
#if SGZ8_TW_BROADCAST == 2
		// one running power (w_k = w_(k-1) w1): four registers fewer than the tree below, three products deeper
		float2 Xk = X1, Yk = Y1;
		apply(1, X1, Y1);
#pragma unroll
		for (int k = 2; k < 8; ++k)
		{
			float2 Xn, Yn;
			product(Xk, Yk, X1, Y1, Xn, Yn);
			Xk = Xn; Yk = Yn;
			apply(k, Xk, Yk);
		}
		return;
#endif
		float2 X2, Y2, X3, Y3, X4, Y4, Xn, Yn;
		product(X1, Y1, X1, Y1, X2, Y2);
		apply(1, X1, Y1);
		apply(2, X2, Y2);
		product(X2, Y2, X1, Y1, X3, Y3);
		apply(3, X3, Y3);
		product(X2, Y2, X2, Y2, X4, Y4);
		apply(4, X4, Y4);
		product(X4, Y4, X1, Y1, Xn, Yn);
		apply(5, Xn, Yn);
		product(X3, Y3, X3, Y3, Xn, Yn);
		apply(6, Xn, Yn);
		product(X4, Y4, X3, Y3, Xn, Yn);
		apply(7, Xn, Yn);
#else
		const float2 w2 = cmul(w1, w1);
		mulTwiddleP(re[perm8(1)], im[perm8(1)], make_float4(w1.x, w1.x, w1.y, w1.y));
		mulTwiddleP(re[perm8(2)], im[perm8(2)], make_float4(w2.x, w2.x, w2.y, w2.y));
		const float2 w3 = cmul(w2, w1);
		mulTwiddleP(re[perm8(3)], im[perm8(3)], make_float4(w3.x, w3.x, w3.y, w3.y));
		const float2 w4 = cmul(w2, w2);
		mulTwiddleP(re[perm8(4)], im[perm8(4)], make_float4(w4.x, w4.x, w4.y, w4.y));
		const float2 w5 = cmul(w4, w1);
		mulTwiddleP(re[perm8(5)], im[perm8(5)], make_float4(w5.x, w5.x, w5.y, w5.y));
		const float2 w6 = cmul(w3, w3);
		mulTwiddleP(re[perm8(6)], im[perm8(6)], make_float4(w6.x, w6.x, w6.y, w6.y));
		const float2 w7 = cmul(w4, w3);
		mulTwiddleP(re[perm8(7)], im[perm8(7)], make_float4(w7.x, w7.x, w7.y, w7.y));
#endif
	}

	// Which bins a thread owns after the last pass: t + T k4.  The two-for-one split pairs bin j with bin N - j, i.e. thread
	// t with thread T - t (slot k4 with slot 7 - k4); this map puts the two partners in lanes l and l ^ 16 of one warp, so
	// that the split can swap them with warp shuffles instead of a shared-memory round trip: lanes 0-15 of warp w hold
	// i = 16 w .. 16 w + 15, lanes 16-31 hold T - i.  The two self-paired threads (t = 0 and t = T / 2) share lane pair 0 / 16
	// of warp 0.  Quarter warps still hold 8 consecutive values of t, which is all the exchange swizzle needs in pass 4.
	template<int T>
	__device__ __forceinline__ int pairedBin(int u)
	{
#if SGZ8_SPLIT_SMEM
		return u;
#endif
		const int i = 16 * (u >> 5) + (u & 15);
		return (u & 16) ? (i == 0 ? T / 2 : T - i) : i;
	}

	// block-wide barrier, or (bar > 0) the named barrier of one 512-thread half of a DUAL CTA
	__device__ __forceinline__ void ctaSync(int bar)
	{
		if (bar == 0) __syncthreads();
		else asm volatile("bar.sync %0, 512;" :: "r"(bar) : "memory");
	}

	// The FFT of one frame once its 8 windowed (E, O) pairs per thread sit in registers: first pass in registers,
	// three in-place radix-8 passes through the exchange buffer, the last radix-2 that joins the even and odd halves.
	// On return re[perm8(k4)] = (Z[k].re, Z[k + M].re), im[...] likewise, for the bins k = t + T k4 (t = pairedBin(u); `wt`
	// points at the W_N^k table, from which W_N^t is fetched just in time for the last radix-2).  Contains the
	// block-wide barriers of the exchanges; the first one also orders whatever the caller still had in xbuf before it.
	template<int K1>
	__device__ __forceinline__ void fusedFftCore(float2 (&re)[8], float2 (&im)[8], float4* xbuf, const float2* twA, const float2* twB, const float2* twC,
		float2 wu, int u, int t, const float2* __restrict__ twN, int bar = 0, int turn = 0)
	{
		constexpr int T = Fused8GeomT<K1>::T;

		if constexpr (K1 <= 8)
		{
			if constexpr (K1 == 8)
			{
				dft8q(re, im);
	#if SGZ8_TW_POWERS
				applyTwiddlePowers(cmul(wu, wu), re, im); // W_M^((64 n2 + v) k1) = (W_M^u)^k1 and W_M^u = (W_N^u)^2
	#else
				{
					const int n2 = u >> 6, v = u & 63;
	#pragma unroll
					for (int k1 = 1; k1 < 8; ++k1)
					{
						const float2 w = cmul(twA[k1 * 8 + n2], twB[k1 * 64 + v]);
						mulTwiddleP(re[perm8(k1)], im[perm8(k1)], make_float4(w.x, w.x, w.y, w.y));
					}
				}
	#endif
			}
			else if constexpr (K1 > 1)
			{
				// G = 8 / K1 radix-K1 butterflies: loaded element e sits at pair index e T + u = n1 (M / K1) + v with
				// n1 = e / G and v = u + T (e % G); output k1 of group g = e % G lands in slot g + G k1 and takes the twiddle
				// (W_M^v)^k1, where W_M^v = W_M^u W_8^g because M / T = 8
				constexpr int G = 8 / K1;
				constexpr float R = 0.70710678118654752440f;
				const float2 wa = cmul(wu, wu); // W_M^u
	#pragma unroll
				for (int g = 0; g < G; ++g)
				{
					float2 w1 = wa;
					if (g == 1) w1 = make_float2(R * (wa.x + wa.y), R * (wa.y - wa.x));   // W_8^1 = (1 - i) / sqrt 2
					if (g == 2) w1 = make_float2(wa.y, -wa.x);                            // W_8^2 = -i
					if (g == 3) w1 = make_float2(R * (wa.y - wa.x), -R * (wa.x + wa.y));  // W_8^3 = -(1 + i) / sqrt 2
					if constexpr (K1 == 4)
					{
						radix4p(re[g], im[g], re[g + 2], im[g + 2], re[g + 4], im[g + 4], re[g + 6], im[g + 6]);
						const float2 w2 = cmul(w1, w1), w3 = cmul(w2, w1);
						mulTwiddleP(re[g + 2], im[g + 2], make_float4(w1.x, w1.x, w1.y, w1.y));
						mulTwiddleP(re[g + 4], im[g + 4], make_float4(w2.x, w2.x, w2.y, w2.y));
						mulTwiddleP(re[g + 6], im[g + 6], make_float4(w3.x, w3.x, w3.y, w3.y));
					}
					else
					{
						const float2 ar = re[g], ai = im[g], br = re[g + 4], bi = im[g + 4];
						re[g] = add2(ar, br); im[g] = add2(ai, bi);
						re[g + 4] = sub2(ar, br); im[g + 4] = sub2(ai, bi);
						mulTwiddleP(re[g + 4], im[g + 4], make_float4(w1.x, w1.x, w1.y, w1.y));
					}
				}
			}
			// (DUAL with turns: the other half of the CTA is through the exchange-heavy part of its frame)
			if (turn) asm volatile("bar.sync %0, 1024;" :: "r"(turn) : "memory");
			ctaSync(bar); // previous frame's csfmag and the staged audio are no longer needed
			if constexpr (K1 == 8)
			{
				const int n2 = u >> 6, n3 = (u >> 3) & 7, n4 = u & 7;
	#pragma unroll
				for (int k1 = 0; k1 < 8; ++k1)
				{
					const float2 r = re[perm8(k1)], i = im[perm8(k1)];
					xbuf[phys8<K1>(k1, n2, n3, n4)] = make_float4(r.x, r.y, i.x, i.y);
				}
			}
			else
			{
				constexpr int G = 8 / K1;
	#pragma unroll
				for (int g = 0; g < G; ++g)
				{
					const int v = u + T * g, n2 = v >> 6, n3 = (v >> 3) & 7, n4 = v & 7;
	#pragma unroll
					for (int k1 = 0; k1 < K1; ++k1)
					{
						const float2 r = re[g + G * k1], i = im[g + G * k1];
						xbuf[phys8<K1>(k1, n2, n3, n4)] = make_float4(r.x, r.y, i.x, i.y);
					}
				}
			}
			ctaSync(bar);
		}
		else
		{
			// K1 = 16 (N = 16384, 1024 threads): the radix-16 first pass is shared by the thread pair (u, u ^ 512).  Loaded
			// element e sits at pair index e T + u = (2 e + h) 512 + v with h = u >> 9, v = u & 511, so thread h holds the
			// n1 of parity h: it takes their DFT8 Y_h[k], the pair swaps them through the exchange buffer, and thread h
			// forms the outputs k1 = k + 8 h:  X[k] = Y_0[k] + W_16^k Y_1[k],  X[k + 8] = Y_0[k] - W_16^k Y_1[k].
			const int h = u >> 9, v = u & 511; // h is uniform over a warp
			dft8q(re, im);
			if (h)
			{
#pragma unroll
				for (int k = 1; k < 8; ++k)
				{
					const float2 w = w32(2 * k);
					mulTwiddleP(re[perm8(k)], im[perm8(k)], make_float4(w.x, w.x, w.y, w.y));
				}
			}
			ctaSync(bar); // previous frame's csfmag and the staged audio are no longer needed
#pragma unroll
			for (int k = 0; k < 8; ++k)
			{
				const float2 r = re[perm8(k)], i = im[perm8(k)];
				xbuf[k * T + u] = make_float4(r.x, r.y, i.x, i.y);
			}
			ctaSync(bar);
#pragma unroll
			for (int k = 0; k < 8; ++k)
			{
				const float4 p = xbuf[k * T + (u ^ 512)];
				const float2 pr = make_float2(p.x, p.y), pi = make_float2(p.z, p.w);
				float2& r = re[perm8(k)]; float2& i = im[perm8(k)];
				if (!h) { r = add2(r, pr); i = add2(i, pi); }
				else { r = sub2(pr, r); i = sub2(pi, i); }
			}
			// pass-1 twiddle (W_M^v)^(k + 8 h), with W_M^v = (W_N^v)^2 and W_N^v = W_N^u W_32^(-h)
			const float2 c1 = w32(1);
			const float2 wv = h ? cmul(wu, make_float2(c1.x, -c1.y)) : wu;
			const float2 w1 = cmul(wv, wv);
			if (h)
			{
				const float2 w2 = cmul(w1, w1), w4 = cmul(w2, w2), w8 = cmul(w4, w4);
#pragma unroll
				for (int k = 0; k < 8; ++k) mulTwiddleP(re[k], im[k], make_float4(w8.x, w8.x, w8.y, w8.y));
			}
			applyTwiddlePowers(w1, re, im);
			ctaSync(bar); // every partner read is done before the buffer is rewritten
			{
				const int n2 = v >> 6, n3 = (v >> 3) & 7, n4 = v & 7;
#pragma unroll
				for (int k = 0; k < 8; ++k)
				{
					const float2 r = re[perm8(k)], i = im[perm8(k)];
					xbuf[phys8<K1>(8 * h + k, n2, n3, n4)] = make_float4(r.x, r.y, i.x, i.y);
				}
			}
			ctaSync(bar);
		}

		SGZ8_TICK(1); // pass 1 written, barrier passed
		// ---- pass 2 (in place): thread (k1 n3 n4), DFT8 over n2, twiddle W_512^(v k2) ----
		{
			const int k1 = u >> 6, n3 = (u >> 3) & 7, n4 = u & 7, v = u & 63;
#pragma unroll
			for (int n2 = 0; n2 < 8; ++n2)
			{
				const float4 x = xbuf[phys8<K1>(k1, n2, n3, n4)];
				re[n2] = make_float2(x.x, x.y); im[n2] = make_float2(x.z, x.w);
			}
			dft8q(re, im);
#if SGZ8_TW_POWERS
			applyTwiddlePowers(twC[64 + v], re, im); // W_512^(v k2) = (W_512^v)^k2
#pragma unroll
			for (int k2 = 0; k2 < 8; ++k2)
			{
				const float2 r = re[perm8(k2)], i = im[perm8(k2)];
				xbuf[phys8<K1>(k1, k2, n3, n4)] = make_float4(r.x, r.y, i.x, i.y);
			}
#else
#pragma unroll
			for (int k2 = 0; k2 < 8; ++k2)
			{
				float2 r = re[perm8(k2)], i = im[perm8(k2)];
				if (k2 > 0)
				{
					const float2 w = twC[k2 * 64 + v];
					mulTwiddleP(r, i, make_float4(w.x, w.x, w.y, w.y));
				}
				xbuf[phys8<K1>(k1, k2, n3, n4)] = make_float4(r.x, r.y, i.x, i.y);
			}
#endif
		}
#if SGZ8_PAIR_BARRIER
		// passes 2 and 3 stay inside the 512-point sub-transform k1 = u >> 6, i.e. inside one pair of warps: only they wait
		if constexpr (K1 <= 8) asm volatile("bar.sync %0, 64;" :: "r"(1 + (u >> 6)) : "memory");
		else ctaSync(bar);
#else
		ctaSync(bar);
#endif

		SGZ8_TICK(2); // pass 2 + barrier
		// ---- pass 3 (in place): thread (k1 k2 n4), DFT8 over n3, twiddle W_64^(n4 k3) ----
		{
			const int k1 = u >> 6, k2 = (u >> 3) & 7, n4 = u & 7;
#pragma unroll
			for (int n3 = 0; n3 < 8; ++n3)
			{
				const float4 x = xbuf[phys8<K1>(k1, k2, n3, n4)];
				re[n3] = make_float2(x.x, x.y); im[n3] = make_float2(x.z, x.w);
			}
			dft8q(re, im);
#if SGZ8_TW_POWERS >= 2
			applyTwiddlePowers(twA[8 + n4], re, im); // W_64^(n4 k3) = (W_64^n4)^k3
#pragma unroll
			for (int k3 = 0; k3 < 8; ++k3)
			{
				const float2 r = re[perm8(k3)], i = im[perm8(k3)];
				xbuf[phys8<K1>(k1, k2, k3, n4)] = make_float4(r.x, r.y, i.x, i.y);
			}
#else
#pragma unroll
			for (int k3 = 0; k3 < 8; ++k3)
			{
				float2 r = re[perm8(k3)], i = im[perm8(k3)];
				if (k3 > 0)
				{
					const float2 w = twA[k3 * 8 + n4];
					mulTwiddleP(r, i, make_float4(w.x, w.x, w.y, w.y));
				}
				xbuf[phys8<K1>(k1, k2, k3, n4)] = make_float4(r.x, r.y, i.x, i.y);
			}
#endif
		}
		ctaSync(bar);

		SGZ8_TICK(3); // pass 3 + barrier
		// ---- pass 4: thread t = k1 + K1 k2 + 8 K1 k3, DFT8 over n4; owns bins k = t + T k4 ----
#if SGZ8_SPLIT_SMEM
		const float2 wt = wu;
#else
		const float2 wt = ldg_f2(twN + t); // W_N^t, in flight behind the exchange reads
#endif
		{
			const int k1 = t & (K1 - 1), k2 = (t / K1) & 7, k3 = t / (8 * K1);
#pragma unroll
			for (int n4 = 0; n4 < 8; ++n4)
			{
				const float4 x = xbuf[phys8<K1>(k1, k2, k3, n4)];
				re[n4] = make_float2(x.x, x.y); im[n4] = make_float2(x.z, x.w);
			}
		}
		dft8q(re, im);
		// last radix-2: re[slot] = (Zp.re, Zm.re), im[slot] = (Zp.im, Zm.im); W_N^(t + T k4) = W_N^t W_16^k4
#pragma unroll
		for (int k4 = 0; k4 < 8; ++k4)
		{
			const int slot = perm8(k4);
			const float2 o = make_float2(re[slot].y, im[slot].y), e = make_float2(re[slot].x, im[slot].x);
			const float2 wo = cmul(cmul(o, w32(2 * k4)), wt);
			re[slot] = make_float2(e.x + wo.x, e.x - wo.x);
			im[slot] = make_float2(e.y + wo.y, e.y - wo.y);
		}
	}


	// doTransform for the staged calls at N = 2048 / 4096 / 8192: the same register-resident FFT as the fused kernel,
	// in place on audioMemory (csf, [stream][N + 1], element N untouched), one CTA per stream.  The complex input
	// z[n] is read as the (even, odd) pairs the core works on; bins come out in natural order.
	template<int K1>
	struct Fft8Smem { static constexpr size_t bytes = Fused8GeomT<K1>::twOffset + Fused8GeomT<K1>::twBytes; };

	// Sub-transform q = blockIdx.y of `subs` interleaved ones: input element n at src[q + subs n], bin k to dst[q + subs k]
	// (subs = 1: the whole transform, in place when src == dst - every thread has read its inputs before the first
	// barrier inside the core, and stores only after the last).  With subs > 1 this finishes a transform of
	// subs * 2 * M points whose first log2(subs) Stockham passes ran through global memory (k_fft_global_pass leaves
	// exactly this layout).  The tables in `t` are those of the 2 M-point transform.
	template<int K1>
	__global__ void __launch_bounds__(64 * K1) k_fft8(DeviceTables t, const float2* src, uint64_t srcStreamStride, float2* dst, uint64_t dstStreamStride, uint32_t subs)
	{
		typedef Fused8GeomT<K1> Gm;
		constexpr int T = Gm::T, M = Gm::M;
		extern __shared__ __align__(16) unsigned char smemRaw[];
		float4* xbuf = reinterpret_cast<float4*>(smemRaw);
		float2* twA = reinterpret_cast<float2*>(smemRaw + Gm::twOffset);
#if SGZ8_TW_POWERS >= 1
		float2* twB = nullptr;
		float2* twC = twA;                                                // twC[64 + v] is W_512^v
#else
		float2* twB = twA + 64;
		float2* twC = twB + 512;
#endif
		const int u = threadIdx.x;
		const float2* z = src + uint64_t(blockIdx.x) * srcStreamStride + blockIdx.y;
		float2* y = dst + uint64_t(blockIdx.x) * dstStreamStride + blockIdx.y;
#if SGZ8_TW_POWERS >= 1
		for (int i = u; i < Gm::twEntries; i += T) twA[i] = t.tw8[i < 64 ? i : 64 + 512 + i]; // A, then row 1 of C
#else
		for (int i = u; i < Gm::twEntries; i += T) twA[i] = t.tw8[i];
#endif
		const float2 wu = t.tw3[u];
		float2 re[8], im[8];
#pragma unroll
		for (int e = 0; e < 8; ++e)
		{
			const uint64_t p = uint64_t(e * T + u) * 2;
			const float2 ev = z[p * subs], od = z[(p + 1) * subs];
			re[e] = make_float2(ev.x, od.x); im[e] = make_float2(ev.y, od.y);
		}
		__syncthreads(); // the twiddle tables are in place
		const int tb = pairedBin<T>(u);
		fusedFftCore<K1>(re, im, xbuf, twA, twB, twC, wu, u, tb, t.tw3);
#pragma unroll
		for (int k4 = 0; k4 < 8; ++k4)
		{
			const int slot = perm8(k4);
			const uint64_t k = uint64_t(tb + T * k4);
			y[k * subs] = make_float2(re[slot].x, im[slot].x);
			y[(k + M) * subs] = make_float2(re[slot].y, im[slot].y);
		}
	}

	__device__ __forceinline__ void prefetchL2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

	// One stored result against the decayed carry: raised where the carry dominates (fmaxf, as the fix-up always did).
	// Returns whether the carry can still matter at LATER frames of the part.  It cannot once scale(carry) is strictly
	// below the stored scale(s): the scale is monotone, so carry < s, and s_(f+1) >= fl(s pole) >= fl(carry pole) keeps it
	// so for good; nor once the scaled carry sits on the clip value, the least value the scale has.  (Equal scaled values
	// do not count: two different states may share one, and part again a frame later.)
	template<bool ASK>
	__device__ __forceinline__ bool seamRaise(float& res, float carried, float sm, const DeviceTables& t)
	{
		const float c = fastLogScale(sm, carried, 1.f, t.deltaYLn2, t.lowerClip);
		const bool live = ASK && sm * carried > 0.f && !(c < res);
		res = fmaxf(res, c);
		return live;
	}
	// whether a raise changed any bit of a result word (a NaN that stays a NaN is no change)
	__device__ __forceinline__ bool differs(float2 a, float2 b) { return ((__float_as_uint(a.x) ^ __float_as_uint(b.x)) | (__float_as_uint(a.y) ^ __float_as_uint(b.y))) != 0u; }
	__device__ __forceinline__ bool differs(float4 a, float4 b)
	{
		return ((__float_as_uint(a.x) ^ __float_as_uint(b.x)) | (__float_as_uint(a.y) ^ __float_as_uint(b.y)) | (__float_as_uint(a.z) ^ __float_as_uint(b.z)) |
			(__float_as_uint(a.w) ^ __float_as_uint(b.w))) != 0u;
	}

	// ---- Phase configuration (MATRIX == 3): the display works on the complex split spectra themselves ----
	// csf as the reference's magnitude pass sees it once the `np` lowest indices and their mirrors have been replaced by
	// their magnitudes (the lazy in-place normalisation of TransformDSP.inl:709-727)
	__device__ __forceinline__ float2 phaseSeen(float2 v, int i, int N, int np)
	{
		return (i < np || i > N - np) ? make_float2(sqrtf(v.x * v.x + v.y * v.y), 0.f) : v;
	}
	__device__ __forceinline__ float2 phaseEntry(const float2* csf, int i, int N) { return (i >= 0 && i <= N) ? csf[i] : make_float2(0.f, 0.f); }
	__device__ __forceinline__ float mag2(float2 v) { return sqrtf(v.x * v.x + v.y * v.y); }

	template<int K1, bool STEREO, int BINPOL, int MATRIX, bool DUAL = false>
#ifndef SGZ8_MINB8
#define SGZ8_MINB8 2 // resident CTAs per SM the 8192-point kernel is compiled for (1: twice the registers per thread, experiment)
#endif
	__global__ void __launch_bounds__(DUAL ? 128 * K1 : 64 * K1, DUAL ? 1 : K1 == 16 ? 1 : K1 == 8 ? SGZ8_MINB8 : K1 == 4 ? 3 : K1 == 2 ? 4 : 5)
	k_fused8(DeviceTables t, FusedArgs a)
	{
		typedef Fused8GeomT<K1> Gm;
		constexpr int T = Gm::T, M = Gm::M, N = Gm::N;
		extern __shared__ __align__(16) unsigned char smemAll[];
		// DUAL: one CTA of 2 T threads runs two frames side by side, each half with its own copy of the whole shared-memory
		// layout, its own range of the launch and its own named barrier - two CTAs in one, which can take turns (below)
		const int half = DUAL ? int(threadIdx.x) / T : 0;
		const int bar = DUAL ? 1 + half : 0;
		unsigned char* const smemRaw = smemAll + (DUAL ? size_t(half) * ((Gm::smemBytes(t.P) + 127) & ~size_t(127)) : 0);
		const uint32_t cta = DUAL ? 2 * blockIdx.x + half : blockIdx.x, nCtas = DUAL ? 2 * gridDim.x : gridDim.x; // this half's range of the launch
		float4* xbuf = reinterpret_cast<float4*>(smemRaw);
		float4* stage = reinterpret_cast<float4*>(smemRaw);               // [4][T] next frame's newest half
		float* csfmag = reinterpret_cast<float*>(smemRaw + size_t(M) * 8); // [N+1], upper half of xbuf (+4 B)
		float2* csfC = reinterpret_cast<float2*>(smemRaw);                 // Phase configuration: [N+1] complex csf = the whole xbuf (+8 B of the pad)
		float2* z0c = reinterpret_cast<float2*>(smemRaw + Gm::xbufBytes + 8);  // Complex configuration: the halved DC bin (in the pad behind csfmag[N])
		float2* twA = reinterpret_cast<float2*>(smemRaw + Gm::twOffset);  // [8][8]   W_64^(a b)
#if SGZ8_TW_POWERS >= 1
		float2* twB = nullptr;
		float2* twC = twA;                                                // twC[64 + v] is W_512^v
#else
		float2* twB = twA + 64;                                           // [8][64]  W_M^(v k1)
		float2* twC = twB + 512;                                          // [8][64]  W_512^(v k2)
#endif
		float2* st = reinterpret_cast<float2*>(smemRaw + Gm::stateOffset); // [2][P]

		const int u = DUAL ? int(threadIdx.x) - half * T : int(threadIdx.x);
		const int tb = pairedBin<T>(u); // this thread owns bins tb + T k4 after the FFT
		const int P = int(t.P);
		const int conf = t.conf;
		const int linLimit = int(t.linLimit);
		uint32_t* mapPacked = reinterpret_cast<uint32_t*>(smemRaw + Gm::mapOffset(t.P));
		float* mapSlope = reinterpret_cast<float*>(mapPacked + P);
		int* mapBaseR = reinterpret_cast<int*>(mapSlope + P);

		for (int i = u; i < P; i += T)
		{
			const int4 bm = t.binmap[i];
			mapPacked[i] = packBinMap(bm.x, bm.y, bm.z);
			mapSlope[i] = t.slope[i] * t.minFracRecip;
			mapBaseR[i] = t.baseR[i];
		}
#if SGZ8_TW_POWERS >= 1
		for (int i = u; i < Gm::twEntries; i += T) twA[i] = t.tw8[i < 64 ? i : 64 + 512 + i]; // A, then row 1 of C
#else
		for (int i = u; i < Gm::twEntries; i += T) twA[i] = t.tw8[i];
#endif
		const float2 wu = t.tw3[u]; // W_N^u: pass 1 uses its square W_M^u
		// DUAL with turns: [0] frames this half did in its earlier parts, [1] frames both halves have (the shorter range): for that many
		// frames the halves alternate - half 1 starts the exchange-heavy part of frame i (passes, split, magnitudes) when
		// half 0 is through it, half 0 that of frame i + 1 when half 1 is through frame i - so that one half's shared-
		// memory traffic runs against the other's display points instead of against its exchanges
		volatile uint32_t* const turns = reinterpret_cast<volatile uint32_t*>(smemRaw + Gm::partOffset(t.P) + 72);
		if (DUAL && SGZ8_DUAL_TURNS && u == 0)
		{
			const uint64_t items = uint64_t(a.numStreams) * a.nFrames;
			const uint64_t mine = items * (cta + 1) / nCtas - items * cta / nCtas, other = items * ((cta ^ 1u) + 1) / nCtas - items * (cta ^ 1u) / nCtas;
			turns[0] = 0; turns[1] = uint32_t(min(mine, other));
		}
		ctaSync(bar);

#ifdef SGZ8_STAGGER
		// A/B switch: the two CTAs of an SM start half a frame apart instead of in lock step
		if (((blockIdx.x & 1u) ^ (blockIdx.x >= gridDim.x / 2 ? 1u : 0u)) != 0u)
		{
			const long long t0 = clock64();
			while (clock64() - t0 < SGZ8_STAGGER) { }
		}
#endif
		// ---- this CTA's share of the launch: items [i0, i1) of the stream-major frame list, cut into per-stream parts ----
		// (planned one part at a time into shared memory by planFusedPart, so that nothing of the bookkeeping stays in
		// registers across the frame loop, which has none to spare)
		FusedPart* part = reinterpret_cast<FusedPart*>(smemRaw + Gm::partOffset(t.P));
		const int W = int(t.W);
#if SGZ8_BULK_STAGE
		uint32_t stagePhase = 0; // parity of the staging mbarrier's current phase (every copy asked for is waited for, so it carries over the parts)
		if (u == 0) mbarInit(smemRaw + Gm::partOffset(t.P) + 64, 1);
		// (the barrier at the top of the part loop publishes it)
#endif

		for (uint32_t q = 0;; ++q)
		{
		ctaSync(bar); // the previous part's plan is no longer read (and its state has been written out)
#ifdef SGZ8_NO_PARTS // A/B: one CTA per stream, nothing planned
		if (u == 0)
		{
			part->nParts = 1; part->fBegin = 0; part->fEnd = a.nFrames; part->stream = int(cta + a.firstStream); part->publish = 0;
			part->src = a.states + uint64_t(part->stream) * 2 * P; part->dst = a.states + uint64_t(part->stream) * 2 * P;
			part->audio = a.audio + uint64_t(part->stream) * a.streamStride; part->vec4 = 1;
		}
#else
		if (u == 0)
		{
			if (DUAL && SGZ8_DUAL_TURNS && q > 0) turns[0] += part->fEnd - part->fBegin; // (the part just finished; nobody reads between the two barriers)
			planFusedPart(a, q, t.P, W == N, part, cta, nCtas);
		}
#endif
		ctaSync(bar);
#ifdef SGZ8_CTA_CLOCK // diagnostic build only: when each CTA of the launch began and ended (globaltimer, ns) and on which SM
		if (q == 0 && u == 0) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ctaClock0() = g; }
		if (q >= part->nParts && u == 0)
		{
			unsigned long long g; unsigned smid;
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
			asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
			printf("cta %u sm %u begin %llu end %llu\n", cta, smid, ctaClock0(), g);
		}
#endif
#ifdef SGZ8_PHASE_CLOCK
		if (q >= part->nParts && u == 0 && blockIdx.x == 0)
		{
			const long long* b = phaseClockBuffer();
			printf("phase clocks over %lld frames (cycles per frame): load+window %lld | pass1+exchange %lld | pass2 %lld | pass3 %lld | pass4+radix2 %lld | split+magnitudes %lld | rows+barrier+stage %lld | display %lld\n",
				b[17], b[0] / b[17], b[1] / b[17], b[2] / b[17], b[3] / b[17], b[4] / b[17], b[5] / b[17], b[6] / b[17], b[7] / b[17]);
		}
#endif
		if (q >= part->nParts) break;
		// (the frame loop reads the plan from shared memory where it needs it, like the kernel arguments from constant memory:
		// its registers are all spoken for)
		const volatile FusedPart* const plan = part;
		{
			const float2* stateSrc = part->src;
			if (stateSrc) { for (int i = u; i < 2 * P; i += T) st[i] = __ldcg(stateSrc + i); }
			else { for (int i = u; i < 2 * P; i += T) st[i] = make_float2(0.f, 0.f); }
		}
		// (the first block-wide barrier of the frame orders these writes before any display point reads them)

		bool staged = false;

#ifdef SGZ8_PHASE_CLOCK
		if (u == 0 && blockIdx.x == 0 && q == 0) { for (int i = 0; i < 18; ++i) phaseClockBuffer()[i] = 0; }
		SGZ8_TICK(-1);
#endif
		for (uint32_t f = plan->fBegin; f < plan->fEnd; ++f)
		{
			float2 re[8], im[8];
			const uint64_t start = uint64_t(f) * a.hop;
			const float* const sbase = plan->audio;
			const bool vec4 = plan->vec4 != 0;

			// ---- pass 1: load (all loads first), window + channel matrix, DFT8 over n1 ----
			{
				float2 wv[8];
#if SGZ8_WINDOW == 0
#pragma unroll
				for (int n1 = 0; n1 < 8; ++n1) wv[n1] = ldg_f2(reinterpret_cast<const float2*>(t.window) + n1 * T + u);
#endif
				float4 rv[8];
				if (vec4)
				{
					const float4* src = reinterpret_cast<const float4*>(sbase + 2 * start) + u;
#pragma unroll
					for (int n1 = 0; n1 < 4; ++n1) rv[n1] = ldg_stream_f4(src + n1 * T);
					if (staged)
					{
#if SGZ8_BULK_STAGE
						mbarWaitParity(smemRaw + Gm::partOffset(t.P) + 64, stagePhase);
						stagePhase ^= 1u;
#else
						cpAsyncWaitAll();
#endif
#pragma unroll
						for (int n1 = 4; n1 < 8; ++n1) rv[n1] = stage[(n1 - 4) * T + u];
					}
					else
					{
#pragma unroll
						for (int n1 = 4; n1 < 8; ++n1) rv[n1] = ldg_stream_f4(src + n1 * T);
					}
				}
				else
				{
#pragma unroll
					for (int n1 = 0; n1 < 8; ++n1)
					{
						const int n = 2 * (n1 * T + u);
						float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
						if (n < W) fetchLR(sbase, a.layout, a.nSamples, start + n, v.x, v.y);
						if (n + 1 < W) fetchLR(sbase, a.layout, a.nSamples, start + n + 1, v.z, v.w);
						rv[n1] = v;
					}
				}
#if SGZ8_WINDOW == 2
				if (t.winPolyOk)
				{
					// cosine-sum window over the whole transform: w = sum_k (-1)^k a_k cos(k theta) is a polynomial p in
					// c = cos(theta), and theta of sample 2 (n1 T + u) + parity is 2 arg(conj W_N^u) + n1 pi / 4 (+ 2 pi / N):
					// nothing is read, eight multiples of 45 degrees share the even / odd halves of p
					constexpr float R = 0.70710678118654752440f;
					const float zx = wu.x * wu.x - wu.y * wu.y, zy = -2.f * wu.x * wu.y;            // e^(i theta) of the even sample, n1 = 0
					const float ox = zx * t.winStep.x - zy * t.winStep.y, oy = zx * t.winStep.y + zy * t.winStep.x; // ... of the odd one
					const float2 c = make_float2(zx, ox), sn = make_float2(zy, oy);
					const float2 arg[4] = { c, mul2(sub2(c, sn), bc2(R)), sub2(bc2(0.f), sn), mul2(add2(c, sn), bc2(-R)) }; // cos at n1 = 0..3
#pragma unroll
					for (int q4 = 0; q4 < 4; ++q4)
					{
						const float2 x = arg[q4], x2 = mul2(x, x);
						const float2 ev = fma2(x2, fma2(x2, bc2(t.winPoly[4]), bc2(t.winPoly[2])), bc2(t.winPoly[0]));
						const float2 od = mul2(x, fma2(x2, bc2(t.winPoly[3]), bc2(t.winPoly[1])));
						wv[q4] = add2(ev, od); wv[q4 + 4] = sub2(ev, od); // cos(theta + pi) = -cos(theta)
					}
				}
				else
#endif
#if SGZ8_WINDOW >= 1
				{
#pragma unroll
					for (int n1 = 0; n1 < 8; ++n1) wv[n1] = ldg_f2(reinterpret_cast<const float2*>(t.window) + n1 * T + u);
				}
#endif
#pragma unroll
				for (int n1 = 0; n1 < 8; ++n1)
				{
					const float4 v = rv[n1];
					const float2 w = wv[n1];
					if constexpr (MATRIX != 1)
					{
						re[n1] = make_float2(v.x * w.x, v.z * w.y);
						im[n1] = make_float2(v.y * w.x, v.w * w.y);
					}
					else
					{
						const float2 e = channelMatrix(conf, v.x, v.y, w.x), o = channelMatrix(conf, v.z, v.w, w.y);
						re[n1] = make_float2(e.x, o.x); im[n1] = make_float2(e.y, o.y);
					}
				}
			}
			SGZ8_TICK(0); // audio + window loaded, multiplied
			int turn = 0;
			if (DUAL && SGZ8_DUAL_TURNS)
			{
				const uint32_t fi = turns[0] + (f - plan->fBegin), both = turns[1];
				if (half == 0 ? (fi > 0 && fi < both) : (fi < both)) turn = half == 0 ? 4 : 3;
			}
			fusedFftCore<K1>(re, im, xbuf, twA, twB, twC, wu, u, tb, t.tw3, bar, turn);
			ctaSync(bar); // every pass-4 read of xbuf is done before csfmag and the staged audio overwrite it
			SGZ8_TICK(4); // pass 4, last radix-2, barrier

#if !SGZ8_STAGE_LATE && !SGZ8_SPLIT_SMEM
			// newest half of the next frame -> lower half of the exchange buffer, in flight behind the split and the display
			// (Phase keeps the whole buffer for its complex spectra)
			staged = false;
			if (MATRIX != 3 && vec4 && f + 1 < plan->fEnd)
			{
#if SGZ8_BULK_STAGE
				// (the barrier above ordered every generic-proxy access of the lower half before this thread; the fence orders them
				// before the async-proxy write)
				if (u == 0) bulkStage(stage, reinterpret_cast<const float4*>(sbase + 2 * (start + a.hop)) + 4 * T, 4 * T * 16, smemRaw + Gm::partOffset(t.P) + 64);
#else
				const float4* src = reinterpret_cast<const float4*>(sbase + 2 * (start + a.hop)) + u;
#pragma unroll
				for (int n1 = 4; n1 < 8; ++n1) cpAsync16(stage + (n1 - 4) * T + u, src + n1 * T);
				cpAsyncCommit();
#endif
				staged = true;
			}
#endif

			if constexpr (MATRIX == 3)
			{
				// Phase keeps the split spectra complex: A (channel 1) at csfC[j], B (channel 2) at csfC[N - j], the whole
				// exchange buffer (TransformDSP.inl:649-652 + separateTransformsIPL).  Z[j + M] is parked at slot M + j first;
				// the partner Z[N - j] of bin j then sits at slot N - j, which is also where B_j goes - every upper slot is
				// read and rewritten by one thread only, and the lower half (the A's) holds nothing that is still needed.
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4) csfC[M + tb + T * k4] = make_float2(re[perm8(k4)].y, im[perm8(k4)].y);
				ctaSync(bar);
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4)
				{
					const int slot = perm8(k4);
					const int j = tb + T * k4;
					const float2 zj = make_float2(re[slot].x, im[slot].x);
					if (k4 == 0 && j == 0)
					{
						const float2 ny = make_float2(re[slot].y, im[slot].y); // Z[M]
						csfC[N] = make_float2(zj.y * 0.5f, 0.f);
						csfC[0] = make_float2(zj.x * 0.5f, 0.f);
						csfC[M] = make_float2(ny.x * 0.5f, ny.y * 0.5f);
						continue;
					}
					const float2 zn = csfC[N - j];
					float2 A = make_float2((zj.x + zn.x) * 0.5f, (zj.y - zn.y) * 0.5f);
					const float2 D = make_float2((zj.x - zn.x) * 0.5f, (zj.y + zn.y) * 0.5f);
					if (k4 == 7 && j == M - 1) { A.x *= 0.5f; A.y *= 0.5f; } // csf[(N >> 1) - 1] *= 0.5
					csfC[j] = A;
					csfC[N - j] = make_float2(D.y, -D.x);
				}
			}
			else 			if constexpr (STEREO)
			{
#if SGZ8_SPLIT_SMEM
				float2* zm = reinterpret_cast<float2*>(smemRaw);
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4) zm[u + T * k4] = make_float2(re[perm8(k4)].y, im[perm8(k4)].y);
				ctaSync(bar);
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4)
				{
					const int slot = perm8(k4);
					const int j = u + T * k4;
					const float2 zj = make_float2(re[slot].x, im[slot].x);
					if (k4 == 0 && j == 0)
					{
						const float2 ny = make_float2(re[slot].y, im[slot].y); // Z[M] = Zm[0]
						const float h2 = cnorm(ny);
						csfmag[0] = zj.x;
						csfmag[N] = zj.y;
						csfmag[M] = (M < linLimit) ? approxSqrt(h2) : h2;
						continue;
					}
					const float2 zn = zm[M - j]; // Z[N - j]
					const float2 S = add2(zj, zn), D = sub2(zj, zn);
					const float2 SS = mul2(S, S), DD = mul2(D, D);
					float a2 = SS.x + DD.y, b2 = DD.x + SS.y;
					if (k4 == 7 && j == M - 1) a2 *= 0.25f; // csf[(N >> 1) - 1] *= 0.5, channel 1 only
					if ((k4 == 0 || linLimit > T) && j < linLimit) { a2 = approxSqrt(a2); b2 = approxSqrt(b2); }
					csfmag[j] = a2;
					csfmag[N - j] = b2;
				}
#else
				// Z[N - j] = Zm[M - j] is held by thread T - t in slot 7 - k4: the lane pair (l, l ^ 16) swaps them by shuffle
				// (pairedBin); lane 16 of warp 0 (t = T / 2) is its own partner, and t = 0 is redone below.
				// A = (Z[j] + conj Z[N-j]) / 2, B = (Z[j] - conj Z[N-j]) / 2i.  With S = Z[j] + Z[N-j] and
				// D = Z[j] - Z[N-j]: 2A = (S.x, D.y), 2|B| components = (D.x, S.y).  The halving is folded into the
				// display scale (magScale = invSize / 2), so entries hold 2|.| (or 4|.|^2 beyond linLimit).
				const int partnerLane = (tb == T / 2 || tb == 0) ? (u & 31) : ((u & 31) ^ 16);
				if (tb == 0)
				{
					// thread 0 pairs with itself one slot further (bin T k4 with Zm[M - T k4] = slot 8 - k4): shift its upper halves
					// down one slot so that the loop below is the same for everybody; Z[0] and Z[M] (DC / Nyquist) are redone after it
					const float2 dc = make_float2(re[perm8(0)].y, im[perm8(0)].y);
#pragma unroll
					for (int m = 0; m < 7; ++m) { re[perm8(m)].y = re[perm8(m + 1)].y; im[perm8(m)].y = im[perm8(m + 1)].y; }
					re[perm8(7)].y = dc.x; im[perm8(7)].y = dc.y; // Z[M], read back below
				}
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4)
				{
					const int slot = perm8(k4), pslot = perm8(7 - k4);
					const int j = tb + T * k4;
					const float2 zj = make_float2(re[slot].x, im[slot].x);
					float2 zn; // Z[N - j]
					zn.x = __shfl_sync(0xffffffffu, re[pslot].y, partnerLane);
					zn.y = __shfl_sync(0xffffffffu, im[pslot].y, partnerLane);
					const float2 S = add2(zj, zn), D = sub2(zj, zn);
					const float2 SS = mul2(S, S), DD = mul2(D, D);
					float a2 = SS.x + DD.y, b2 = DD.x + SS.y;
					if (k4 == 7 && j == M - 1) a2 *= 0.25f; // csf[(N >> 1) - 1] *= 0.5, channel 1 only
					if ((k4 == 0 || linLimit > T) && j < linLimit) { a2 = approxSqrt(a2); b2 = approxSqrt(b2); }
					csfmag[j] = a2;
					csfmag[N - j] = b2;
				}
				if (tb == 0)
				{
					// :861-864: csf[N] = Im Z[0] / 2, csf[0] = Re Z[0] / 2, csf[N/2] = packed Nyquists / 2
					const float2 zj = make_float2(re[perm8(0)].x, im[perm8(0)].x);
					const float2 ny = make_float2(re[perm8(7)].y, im[perm8(7)].y); // Z[M] = Zm[0], parked above
					const float h2 = cnorm(ny);
					csfmag[0] = zj.x;
					csfmag[N] = zj.y;
					csfmag[M] = (M < linLimit) ? approxSqrt(h2) : h2;
				}
			#endif
			}
			else if constexpr (MATRIX == 2)
			{
				// Complex configuration (TransformDSP.inl:993-1002): one channel over all N bins, csf[0] halved and left
				// complex, every other bin replaced by its magnitude; csf[N] is the zero pad.  Entry 0 keeps its complex
				// value in `z0c` for the few display points whose interpolation taps reach it.
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4)
				{
					const int slot = perm8(k4);
					const int j = tb + T * k4;
					float2 zp = make_float2(re[slot].x, im[slot].x);
					const float2 zq = make_float2(re[slot].y, im[slot].y);
					if (j == 0) { zp.x *= 0.5f; zp.y *= 0.5f; *z0c = zp; csfmag[N] = 0.f; }
					csfmag[j] = approxSqrt(cnorm(zp));
					csfmag[j + M] = approxSqrt(cnorm(zq));
				}
			}
			else
			{
#pragma unroll
				for (int k4 = 0; k4 < 8; ++k4)
				{
					const int slot = perm8(k4);
					const int j = tb + T * k4;
					float2 zj = make_float2(re[slot].x, im[slot].x);
					if (j == 0)
					{
						zj.x *= 0.5f; zj.y *= 0.5f;
						const float2 ny = make_float2(re[slot].y, im[slot].y);
						const float h2 = 0.25f * cnorm(ny);
						csfmag[M] = (M < linLimit) ? approxSqrt(h2) : h2;
					}
					const float m2 = cnorm(zj);
					csfmag[j] = (j < linLimit) ? approxSqrt(m2) : m2;
				}
			}
			// where this frame's two graphs go (resultRow: always somewhere, so the stores below are unconditional): sixty
			// instructions of 64-bit index arithmetic, done by one thread and read by all after the barrier (the previous frame's
			// readers passed several barriers ago)
			SGZ8_TICK(5); // split exchange + magnitudes written
			if (DUAL && SGZ8_DUAL_TURNS)
			{
				// through the exchange-heavy part: the other half may start its own (non-blocking)
				const uint32_t fi = turns[0] + (f - plan->fBegin), both = turns[1];
				if (half == 0 ? (fi < both) : (fi + 1 < both)) asm volatile("bar.arrive %0, 1024;" :: "r"(half == 0 ? 3 : 4) : "memory");
			}
			if (u == 0) { part->row[0] = resultRow(a, uint64_t(plan->stream), f, 0, uint32_t(P)); part->row[1] = resultRow(a, uint64_t(plan->stream), f, 1, uint32_t(P)); }
			ctaSync(bar);

#if SGZ8_STAGE_LATE || SGZ8_SPLIT_SMEM
			// newest half of the next frame -> lower half of the exchange buffer, in flight behind the split and the display
			// (Phase keeps the whole buffer for its complex spectra)
			staged = false;
			if (MATRIX != 3 && vec4 && f + 1 < plan->fEnd)
			{
#if SGZ8_BULK_STAGE
				// (the barrier above ordered every generic-proxy access of the lower half before this thread; the fence orders them
				// before the async-proxy write)
				if (u == 0) bulkStage(stage, reinterpret_cast<const float4*>(sbase + 2 * (start + a.hop)) + 4 * T, 4 * T * 16, smemRaw + Gm::partOffset(t.P) + 64);
#else
				const float4* src = reinterpret_cast<const float4*>(sbase + 2 * (start + a.hop)) + u;
#pragma unroll
				for (int n1 = 4; n1 < 8; ++n1) cpAsync16(stage + (n1 - 4) * T + u, src + n1 * T);
				cpAsyncCommit();
#endif
				staged = true;
			}
#endif
			// ---- display points: one per thread per step ----
			SGZ8_TICK(6); // barrier before the display, staging issued
			float2* const resRow[2] = { plan->row[0], plan->row[1] };
			const float poles[2] = { t.pole0, t.pole1 };
			const int safeFrom = int(t.safeFrom);
			const float magScale = STEREO ? 0.5f * t.invSize : t.invSize; // the stereo split stores 2 |.|
			for (int x = u; x < P; x += T)
			{
				const uint32_t pkd = mapPacked[x];
				const int first = int(pkd & 0xFFFFu), count = int((pkd >> 16) & 0x7FFFu);
				if constexpr (MATRIX == 3)
				{
					// Phase (TransformDSP.inl:643-853 + :1395-1424): wsp[2x] = invSize (|A| + |B|) from the partly
					// normalised view of csf, wsp[2x+1] = 1 - |A + B| / (|A| + |B|) from the raw one
					const float invSize = t.invSize;
					float mid, cancelOut;
					if (!(pkd >> 31))
					{
						float2 iL, iR, mL, mR;
						if constexpr (BINPOL == SGZ_BIN_NONE)
						{
							iL = csfC[first]; iR = csfC[N - first]; mL = iL; mR = iR;
						}
						else if constexpr (BINPOL == SGZ_BIN_LINEAR)
						{
							const int np = __ldg(t.phaseNormPos + x);
							{
								const float pos = __ldg(t.posL + x); const float fl = floorf(pos); const int k = int(fl); const float fr = pos - fl;
								const float2 a0 = phaseEntry(csfC, k, N), b0 = phaseEntry(csfC, k + 1, N);
								const float2 a1 = phaseSeen(a0, k, N, np), b1 = phaseSeen(b0, k + 1, N, np);
								iL = make_float2(a0.x * (1.f - fr) + b0.x * fr, a0.y * (1.f - fr) + b0.y * fr);
								mL = make_float2(a1.x * (1.f - fr) + b1.x * fr, a1.y * (1.f - fr) + b1.y * fr);
							}
							{
								const float pos = __ldg(t.posR + x); const float fl = floorf(pos); const int k = int(fl); const float fr = pos - fl;
								const float2 a0 = phaseEntry(csfC, k, N), b0 = phaseEntry(csfC, k + 1, N);
								const float2 a1 = phaseSeen(a0, k, N, np), b1 = phaseSeen(b0, k + 1, N, np);
								iR = make_float2(a0.x * (1.f - fr) + b0.x * fr, a0.y * (1.f - fr) + b0.y * fr);
								mR = make_float2(a1.x * (1.f - fr) + b1.x * fr, a1.y * (1.f - fr) + b1.y * fr);
							}
						}
						else
						{
							const int np = __ldg(t.phaseNormPos + x);
							const int bl = first - (kLanczosSize - 1), br = mapBaseR[x];
							const float4* wl4 = t.wL4 + x; // [3][P]: consecutive lanes read consecutive 16-byte words
							const float4* wr4 = t.wR4 + x;
							const float4 w0 = __ldg(wl4), w1 = __ldg(wl4 + P), w2 = __ldg(wl4 + 2 * P);
							const float4 r0 = __ldg(wr4), r1 = __ldg(wr4 + P), r2 = __ldg(wr4 + 2 * P);
							const float w[10] = { w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y };
							const float v[10] = { r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y };
							iL = iR = mL = mR = make_float2(0.f, 0.f);
#pragma unroll
							for (int k = 0; k < kLanczosTaps; ++k)
							{
								const float2 a0 = phaseEntry(csfC, bl + k, N), b0 = phaseEntry(csfC, br + k, N);
								const float2 a1 = phaseSeen(a0, bl + k, N, np), b1 = phaseSeen(b0, br + k, N, np);
								iL.x += a0.x * w[k]; iL.y += a0.y * w[k]; mL.x += a1.x * w[k]; mL.y += a1.y * w[k];
								iR.x += b0.x * v[k]; iR.y += b0.y * v[k]; mR.x += b1.x * v[k]; mR.y += b1.y * v[k];
							}
						}
						const float cancellation = invSize * mag2(make_float2(iL.x + iR.x, iL.y + iR.y));
						const float midRaw = invSize * (mag2(iL) + mag2(iR));
						cancelOut = 1.f - (midRaw > 0.f ? cancellation / midRaw : 0.f);
						mid = invSize * (mag2(mL) + mag2(mR));
					}
					else
					{
						const int np = t.phaseNormFinal;
						int maxBin = 0; float maxValue = 0.f;
						for (int c = 0; c < count; ++c)
						{
							const int off = first + c;
							const float m = fmaxf(cnorm(phaseSeen(csfC[off], off, N, np)), cnorm(phaseSeen(csfC[N - off], N - off, N, np)));
							if (m > maxValue) { maxValue = m; maxBin = off; }
						}
						const float2 l = phaseSeen(csfC[maxBin], maxBin, N, np), r = phaseSeen(csfC[N - maxBin], N - maxBin, N, np);
						const float interference = invSize * mag2(make_float2(l.x + r.x, l.y + r.y));
						mid = invSize * (mag2(l) + mag2(r));
						cancelOut = 1.f - (mid > 0.f ? interference / mid : 0.f);
					}
					const float mag = mid * 0.5f;
					float ph = cancelOut;
					const float sm = mapSlope[x];
					const float phasePoles[2] = { t.phasePole0, t.phasePole1 };
#pragma unroll
					for (int k = 0; k < 2; ++k)
					{
						float2 sv = st[k * P + x];
						sv.x = fmaxf(sv.x * poles[k], mag);
						ph *= mag; // inside the graph loop, as the reference (:1417)
						sv.y = ph + phasePoles[k] * (sv.y - ph);
						st[k * P + x] = sv;
						float2 res;
						res.x = fastLogScale(sm, sv.x, 1.f, t.deltaYLn2, t.lowerClip);
						res.y = fastLogScale(sm, sv.y, 1.f, t.deltaYLn2, t.lowerClip);
						// frame-parallel mode: the smoothing runs over ALL frames afterwards (k_phase_smooth), which needs this
						// frame's filter input rather than the output of a filter that started from nothing
						if (!a.chained) res.y = ph;
						stg_stream_f2(resRow[k] + x, res);
					}
					continue;
				}
				float vL, vR = 0.f;
				if (!(pkd >> 31))
				{
					if constexpr (BINPOL == SGZ_BIN_NONE)
					{
						vL = csfmag[first];
						if constexpr (STEREO) vR = csfmag[N - first];
					}
					else if constexpr (BINPOL == SGZ_BIN_LINEAR)
					{
						{
							const float pos = __ldg(t.posL + x); const float fl = floorf(pos); const int k = int(fl); const float fr = pos - fl;
							vL = csfTap(csfmag, k, N) * (1.f - fr) + csfTap(csfmag, k + 1, N) * fr;
							if constexpr (MATRIX == 2)
								if (k == 0 || k == -1)
								{
									// a tap on the complex DC bin: swap its magnitude for its value, then take the magnitude of the sum
									const float w0 = k == 0 ? 1.f - fr : fr;
									const float2 z0 = *z0c;
									const float vr = vL + w0 * (z0.x - csfmag[0]), vi = w0 * z0.y;
									vL = sqrtf(vr * vr + vi * vi);
								}
						}
						if constexpr (STEREO)
						{
							const float pos = __ldg(t.posR + x); const float fl = floorf(pos); const int k = int(fl); const float fr = pos - fl;
							vR = csfTap(csfmag, k, N) * (1.f - fr) + csfTap(csfmag, k + 1, N) * fr;
						}
					}
					else
					{
						const int bl = first - (kLanczosSize - 1);
#if SGZ8_AOS_WEIGHTS
						const float4* wl4 = reinterpret_cast<const float4*>(t.wL + size_t(x) * kLanczosStride);
						const float4 w0 = __ldg(wl4), w1 = __ldg(wl4 + 1), w2 = __ldg(wl4 + 2);
#else
						const float4* wl4 = t.wL4 + x; // [3][P]: consecutive lanes read consecutive 16-byte words
						const float4 w0 = __ldg(wl4), w1 = __ldg(wl4 + P), w2 = __ldg(wl4 + 2 * P);
#endif
						float4 r0, r1, r2; int br = 0;
						if constexpr (STEREO)
						{
#if SGZ8_AOS_WEIGHTS
							const float4* wr4 = reinterpret_cast<const float4*>(t.wR + size_t(x) * kLanczosStride);
							r0 = __ldg(wr4); r1 = __ldg(wr4 + 1); r2 = __ldg(wr4 + 2);
#else
							const float4* wr4 = t.wR4 + x;
							r0 = __ldg(wr4); r1 = __ldg(wr4 + P); r2 = __ldg(wr4 + 2 * P);
#endif
							br = mapBaseR[x];
						}
						const float w[10] = { w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y };
						vL = 0.f;
						if (x >= safeFrom)
						{
#pragma unroll
							for (int k = 0; k < kLanczosTaps; ++k) vL += csfmag[bl + k] * w[k];
						}
						else
						{
#pragma unroll
							for (int k = 0; k < kLanczosTaps; ++k) vL += csfTap(csfmag, bl + k, N) * w[k];
						}
						if constexpr (MATRIX == 2)
							if (bl <= 0 && bl > -kLanczosTaps)
							{
								// a tap on the complex DC bin (see the linear case)
								float w0 = 0.f;
#pragma unroll
								for (int k = 0; k < kLanczosTaps; ++k) if (bl + k == 0) w0 = w[k];
								const float2 z0 = *z0c;
								const float vr = vL + w0 * (z0.x - csfmag[0]), vi = w0 * z0.y;
								vL = sqrtf(vr * vr + vi * vi);
							}
						if constexpr (STEREO)
						{
							const float v[10] = { r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y };
							if (x >= safeFrom)
							{
#pragma unroll
								for (int k = 0; k < kLanczosTaps; ++k) vR += csfmag[br + k] * v[k];
							}
							else
							{
#pragma unroll
								for (int k = 0; k < kLanczosTaps; ++k) vR += csfTap(csfmag, br + k, N) * v[k];
							}
						}
					}
				}
				else
				{
					float bestL = 0.f, bestR = 0.f;
					if (first >= linLimit)
					{
#if SGZ8_SCAN_ROTATE
						// the lanes of a warp scan runs that lie `count` bins apart: with an even count, lanes 32 / g apart
						// (g = the power of two count shares with the 32 banks) meet in one bank at every step.  A maximum does
						// not care where the scan starts, so those lanes start o bins into their run and wrap around.
						const int g = min(count & -count, 32), o = ((u & 31) * g) >> 5;
						for (int c = o; c < count; ++c)
						{
							bestL = fmaxf(bestL, csfmag[first + c]);
							if constexpr (STEREO) bestR = fmaxf(bestR, csfmag[N - first - c]);
						}
						for (int c = 0; c < o; ++c)
						{
							bestL = fmaxf(bestL, csfmag[first + c]);
							if constexpr (STEREO) bestR = fmaxf(bestR, csfmag[N - first - c]);
						}
#else
						for (int c = 0; c < count; ++c)
						{
							bestL = fmaxf(bestL, csfmag[first + c]);
							if constexpr (STEREO) bestR = fmaxf(bestR, csfmag[N - first - c]);
						}
#endif
					}
					else
					{
						for (int c = 0; c < count; ++c)
						{
							const int off = first + c;
							const float v = csfmag[off];
							bestL = fmaxf(bestL, off < linLimit ? v * v : v);
							if constexpr (STEREO) { const float vr = csfmag[N - off]; bestR = fmaxf(bestR, off < linLimit ? vr * vr : vr); }
						}
					}
					const int fb = first + count - 1;
					const float fv = csfmag[fb];
					const float fbMag = (fb < linLimit || fb > N - linLimit) ? fv : approxSqrt(fv);
					vL = bestL > 0.f ? approxSqrt(bestL) : fbMag;
					if constexpr (STEREO) vR = bestR > 0.f ? approxSqrt(bestR) : fbMag;
				}
				const float lmag = fabsf(magScale * vL);
				const float rmag = fabsf(magScale * vR);
				const float sm = mapSlope[x];
#pragma unroll
				for (int k = 0; k < 2; ++k)
				{
					float2 sv = st[k * P + x];
					// st = max(st * pole, mag); y = d > 0 ? ln(d) dy : clip  ==  max(log2(d) dy ln2, clip)
					// (log2(0) = -inf and NaN both fall to clip, exactly like the reference's `d > 0 ?`)
					sv.x = fmaxf(sv.x * poles[k], lmag);
					float2 res;
					res.x = fastLogScale(sm, sv.x, 1.f, t.deltaYLn2, t.lowerClip);
					if constexpr (STEREO)
					{
						sv.y = fmaxf(sv.y * poles[k], rmag);
						res.y = fastLogScale(sm, sv.y, 1.f, t.deltaYLn2, t.lowerClip);
					}
					else res.y = 0.f;
					st[k * P + x] = sv;
					stg_stream_f2(resRow[k] + x, res);
				}
			}
			SGZ8_TICK(7); // this thread's display points
#ifdef SGZ8_PHASE_CLOCK
			if (u == 0 && blockIdx.x == 0) phaseClockBuffer()[17] += 1;
#endif
		}
		ctaSync(bar); // every display point of the part's last frame has updated its state
		{
			float2* stateDst = part->dst;
			for (int i = u; i < 2 * P; i += T) stateDst[i] = st[i];
			if (part->publish)
			{
				ctaSync(bar);
				if (u == 0) { __threadfence(); stRelease(a.flags + cta, a.epoch); }
			}
		}
		} // parts
	}

	// Second half of the unchained launch mode (fewer streams than CTAs).  Every part of a stream went through k_fused8
	// from a zero state (the part that begins the stream from the stream's real one), so its results already hold the
	// peak hold of its OWN frames.  What is missing is the hold carried in from earlier parts:
	// s_f = max(s_(f-1) pole, m_f) unrolls to max(carry pole^k, local_f), float multiplication by the pole and the dB
	// scale are both monotone, hence
	//     result_f = max(local result_f, scale(carry pole^k))        bit for bit what one CTA per stream computes.
	// k_seam_carry: one thread per display point walks the parts of a stream in order WITHOUT touching the results - it
	// decays the carried state once per frame (the same chain of roundings), leaves in each part's slot the state that
	// part starts from, and merges in the part's own final state.  k_seam_apply: one thread per (part, display point)
	// then raises the stored results of its part where the decaying carry still dominates - fully parallel.
	// Parts are the CTA ranges of k_fused8: cut(c) = floor(c I / G); the part of CTA c that began mid-stream keeps its
	// state in seam[c], the part that began a stream in seam[G + c].
	__device__ __forceinline__ uint64_t seamCut(uint64_t items, uint32_t G, uint32_t c) { return items * c / G; }

	__global__ void __launch_bounds__(128) k_seam_carry(DeviceTables t, float2* __restrict__ seam, float2* __restrict__ states, uint32_t nFrames,
		uint32_t nStreams, uint32_t G, int stereo, uint32_t firstStream)
	{
		const int P = int(t.P);
		const int x = blockIdx.x * blockDim.x + threadIdx.x;
		if (x >= P) return;
		const uint32_t sLocal = blockIdx.y;
		const uint64_t items = uint64_t(nStreams) * nFrames, begin = uint64_t(sLocal) * nFrames, end = begin + nFrames;
		// the CTA whose range holds the stream's first frame
		uint32_t c = uint32_t(begin * G / items);
		while (c + 1 < G && seamCut(items, G, c + 1) <= begin) ++c;
		while (c > 0 && seamCut(items, G, c) > begin) --c;
		const float poles[2] = { t.pole0, t.pole1 };
		const float2* firstSlot = seam + uint64_t(G + c) * 2 * P + x;
		float2 carry[2] = { firstSlot[0], firstSlot[P] }; // final state of the stream's first part = the true state there
		// (a long recording is a chain of hundreds of parts: their final states are fetched eight parts at a time, ahead of
		// the walk, which otherwise is one L2 round trip per part - 200 us for one stream in 296 parts)
		constexpr int kAheadParts = 8;
		for (++c; c < G && seamCut(items, G, c) < end; c += kAheadParts)
		{
			float2 own[kAheadParts][2];
			uint64_t cut[kAheadParts + 1]; // (64-bit divisions: once per part)
			cut[0] = seamCut(items, G, c);
#pragma unroll
			for (int j = 0; j < kAheadParts; ++j)
			{
				cut[j + 1] = c + j < G ? seamCut(items, G, c + j + 1) : end;
				if (c + j < G && cut[j] < end)
				{
					const float2* theirs = seam + uint64_t(c + j) * 2 * P + x;
					own[j][0] = theirs[0]; own[j][1] = theirs[P];
				}
			}
#pragma unroll
			for (int j = 0; j < kAheadParts; ++j)
			{
				if (!(c + j < G && cut[j] < end)) break;
				float2* mine = seam + uint64_t(c + j) * 2 * P + x;
				mine[0] = carry[0]; mine[P] = carry[1]; // what this part starts from
				const uint32_t len = uint32_t((cut[j + 1] < end ? cut[j + 1] : end) - cut[j]);
				for (uint32_t f = 0; f < len; ++f)
				{
#pragma unroll
					for (int k = 0; k < 2; ++k) { carry[k].x *= poles[k]; if (stereo) carry[k].y *= poles[k]; }
				}
#pragma unroll
				for (int k = 0; k < 2; ++k)
				{
					carry[k].x = fmaxf(carry[k].x, own[j][k].x);
					if (stereo) carry[k].y = fmaxf(carry[k].y, own[j][k].y);
				}
			}
		}
		// (.y is only a peak hold in the stereo forms; otherwise it is left as it was - unused, or the Phase smoother's)
#pragma unroll
		for (int k = 0; k < 2; ++k)
		{
			float2* st = states + ((uint64_t(sLocal) + firstStream) * 2 + k) * P + x;
			if (stereo) *st = carry[k]; else st->x = carry[k].x;
		}
	}

	// (gridDim.z slices of a part's frames run side by side: a slice first decays the carry over the frames before it -
	// multiplications only, the same chain of roundings - then walks its own frames four at a time.  One thread walking a
	// whole part is a chain of DRAM round trips as long as the part: 77 us for 55 frames, whatever the GPU had left idle.
	// The walk is bound by its instruction count, not by DRAM: V = 2 points per thread and 16-byte accesses when P is even.)
	// FULL: every frame keeps both graphs, one after the other, so the rows of a slice are 2 P results apart and the walk
	// steps a pointer; the other output modes ask keptRow frame by frame.
	template<int V, bool FULL, bool STEREO>
	__global__ void __launch_bounds__(128) k_seam_apply(DeviceTables t, FusedArgs a, uint32_t G)
	{
		const int P = int(t.P);
		const int x = (blockIdx.x * blockDim.x + threadIdx.x) * V;
		if (x >= P) return;
		const uint32_t c = blockIdx.y, nFrames = a.nFrames;
		const uint64_t items = uint64_t(a.numStreams) * nFrames, i0 = seamCut(items, G, c), i1 = seamCut(items, G, c + 1);
		if (i1 == i0) return;
		const uint32_t sLocal = uint32_t(i0 / nFrames), f0 = uint32_t(i0 - uint64_t(sLocal) * nFrames);
		if (f0 == 0) return; // the range begins a stream: nothing is carried in
		const uint32_t f1 = uint32_t(min(uint64_t(nFrames), i1 - uint64_t(sLocal) * nFrames));
		const uint32_t perSlice = ((f1 - f0 + gridDim.z - 1) / gridDim.z + 3) & ~3u;
		const uint32_t fa = f0 + blockIdx.z * perSlice, fb = min(f1, fa + perSlice);
		if (fa >= f1) return;
		typedef typename std::conditional<V == 2, float4, float2>::type Word; // V points' (.x, .y) results
		float sm[V];
#pragma unroll
		for (int v = 0; v < V; ++v) sm[v] = t.slope[x + v] * t.minFracRecip;
		const float poles[2] = { t.pole0, t.pole1 };
		const float2* mine = a.seam + uint64_t(c) * 2 * P + x;
		float2 carry[2][V];
		auto decay = [&](uint32_t frames)
		{
			for (uint32_t f = 0; f < frames; ++f)
			{
#pragma unroll
				for (int k = 0; k < 2; ++k)
#pragma unroll
					for (int v = 0; v < V; ++v) { carry[k][v].x *= poles[k]; if (STEREO) carry[k][v].y *= poles[k]; }
			}
		};
		auto merge = [&](const float2* own) // a part's final state from zero into the carried one: the true state there
		{
#pragma unroll
			for (int k = 0; k < 2; ++k)
#pragma unroll
				for (int v = 0; v < V; ++v)
				{
					const float2 o = own[k * P + v];
					carry[k][v].x = fmaxf(carry[k][v].x, o.x);
					if (STEREO) carry[k][v].y = fmaxf(carry[k][v].y, o.y);
				}
		};
		const uint64_t stream = uint64_t(sLocal) + a.firstStream;
		if (a.inlineCarry)
		{
			// k_seam_carry's walk, done here when a stream has few parts: from the final state of the part that began the
			// stream through the parts between it and this one
			const uint64_t begin = uint64_t(sLocal) * nFrames;
			uint32_t ch = uint32_t(begin * G / items);
			while (ch + 1 < G && seamCut(items, G, ch + 1) <= begin) ++ch;
			while (ch > 0 && seamCut(items, G, ch) > begin) --ch;
			const float2* head = a.seam + uint64_t(G + ch) * 2 * P + x;
#pragma unroll
			for (int k = 0; k < 2; ++k)
#pragma unroll
				for (int v = 0; v < V; ++v) carry[k][v] = head[k * P + v];
			for (uint32_t cc = ch + 1; cc < c; ++cc)
			{
				decay(uint32_t(seamCut(items, G, cc + 1) - seamCut(items, G, cc)));
				merge(a.seam + uint64_t(cc) * 2 * P + x);
			}
			if (f1 == nFrames && blockIdx.z == 0)
			{
				// this part ends the stream: its true final state is the stream's (.y is only a peak hold in the stereo forms;
				// otherwise it is left as it was - unused, or the Phase smoother's)
				float2 keep[2][V];
#pragma unroll
				for (int k = 0; k < 2; ++k)
#pragma unroll
					for (int v = 0; v < V; ++v) keep[k][v] = carry[k][v];
				decay(f1 - f0);
				merge(mine);
#pragma unroll
				for (int k = 0; k < 2; ++k)
#pragma unroll
					for (int v = 0; v < V; ++v)
					{
						float2* st = a.states + (stream * 2 + k) * P + x + v;
						if (STEREO) *st = carry[k][v]; else st->x = carry[k][v].x;
						carry[k][v] = keep[k][v];
					}
			}
		}
		else
		{
#pragma unroll
			for (int k = 0; k < 2; ++k)
#pragma unroll
				for (int v = 0; v < V; ++v) carry[k][v] = mine[k * P + v];
		}
		decay(fa - f0);
		// one frame's two rows against the carry; with `ask`, returns whether the carry can still reach a later frame
		// (seamRaise) - only the last frame of a trip asks
		auto frame = [&](Word* const row[2], Word res[2], auto ask) -> bool
		{
			constexpr bool kAsk = decltype(ask)::value;
			bool live = false;
#pragma unroll
			for (int k = 0; k < 2; ++k)
			{
				Word was = res[k];
				float* val = reinterpret_cast<float*>(&res[k]);
#pragma unroll
				for (int v = 0; v < V; ++v)
				{
					carry[k][v].x *= poles[k];
					if (STEREO) carry[k][v].y *= poles[k];
					if (!row[k]) { if (kAsk) live = live || carry[k][v].x > 0.f || (STEREO && carry[k][v].y > 0.f); continue; }
					live = seamRaise<kAsk>(val[2 * v], carry[k][v].x, sm[v], t) || live;
					if (STEREO) live = seamRaise<kAsk>(val[2 * v + 1], carry[k][v].y, sm[v], t) || live;
				}
				if (row[k] && differs(was, res[k])) *row[k] = res[k];
			}
			return live;
		};
		const std::true_type asking; const std::false_type quiet;
		if constexpr (FULL)
		{
			const size_t rowStep = size_t(P) / V; // Words between one graph and the next
			Word* at = reinterpret_cast<Word*>(a.results + ((stream * nFrames + fa) * 2) * P + x);
			uint32_t left = fb - fa;
			for (; left >= 4; left -= 4, at += 8 * rowStep)
			{
				Word res[4][2];
#pragma unroll
				for (int j = 0; j < 4; ++j) { res[j][0] = __ldcg(at + (2 * j) * rowStep); res[j][1] = __ldcg(at + (2 * j + 1) * rowStep); }
				bool live = false;
#pragma unroll
				for (int j = 0; j < 4; ++j)
				{
					Word* const row[2] = { at + (2 * j) * rowStep, at + (2 * j + 1) * rowStep };
					if (j < 3) frame(row, res[j], quiet); else live = frame(row, res[j], asking);
				}
				if (!live) return;
			}
			for (; left > 0; --left, at += 2 * rowStep)
			{
				Word* const row[2] = { at, at + rowStep };
				Word res[2] = { __ldcg(row[0]), __ldcg(row[1]) };
				if (!frame(row, res, asking)) return;
			}
		}
		else for (uint32_t f = fa; f < fb; f += 4)
		{
			Word* row[4][2]; Word res[4][2];
#pragma unroll
			for (int j = 0; j < 4; ++j)
#pragma unroll
				for (int k = 0; k < 2; ++k)
				{
					float2* r = f + j < fb ? keptRow(a, stream, f + j, k, uint32_t(P)) : nullptr;
					row[j][k] = r ? reinterpret_cast<Word*>(r + x) : nullptr;
					if (r) res[j][k] = __ldcg(row[j][k]);
				}
			bool live = false;
#pragma unroll
			for (int j = 0; j < 4; ++j)
			{
				if (f + j >= fb) break;
				live = frame(row[j], res[j], asking);
			}
			if (!live) break;
		}
	}

	// Frame-parallel mode, Phase configuration: the second state of a display point is a one-pole smoother of
	// ph = cancellation x magnitude (TransformDSP.inl:1417-1419), not a maximum, so it cannot be patched up run by run.
	// The runs leave ph in the .y slot of both graphs and this kernel, one thread per display point, filters it over all
	// frames in order with the reference's expression and writes the dB-scaled value - the same roundings as one CTA per
	// stream.  Four frames are fetched per trip: the walk is latency-bound and the loads do not depend on the filter.
	__global__ void __launch_bounds__(128) k_phase_smooth(DeviceTables t, float2* __restrict__ results, float2* __restrict__ states, uint32_t nFrames,
		uint32_t firstStream)
	{
		const int P = int(t.P);
		const int x = blockIdx.x * blockDim.x + threadIdx.x;
		if (x >= P) return;
		const uint64_t s = blockIdx.y + firstStream;
		const float sm = t.slope[x] * t.minFracRecip;
		const float phasePoles[2] = { t.phasePole0, t.phasePole1 };
		float y[2] = { states[(s * 2 + 0) * P + x].y, states[(s * 2 + 1) * P + x].y };
		float2* frame = results + (s * nFrames * 2) * P + x;
		for (uint32_t f = 0; f < nFrames; f += 4, frame += 8 * P)
		{
			float2 res[4][2];
#pragma unroll
			for (int j = 0; j < 4; ++j)
				if (f + j < nFrames) { res[j][0] = frame[uint64_t(j) * 2 * P]; res[j][1] = frame[uint64_t(j) * 2 * P + P]; }
#pragma unroll
			for (int j = 0; j < 4; ++j)
			{
				if (f + j >= nFrames) break;
#pragma unroll
				for (int k = 0; k < 2; ++k)
				{
					const float ph = res[j][k].y;
					y[k] = ph + phasePoles[k] * (y[k] - ph);
					res[j][k].y = fastLogScale(sm, y[k], 1.f, t.deltaYLn2, t.lowerClip);
					frame[uint64_t(j) * 2 * P + k * P] = res[j][k];
				}
			}
		}
		states[(s * 2 + 0) * P + x].y = y[0];
		states[(s * 2 + 1) * P + x].y = y[1];
	}
