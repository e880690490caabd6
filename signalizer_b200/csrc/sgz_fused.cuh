	// ------------------------------------------------------------------------------------------
	// fused kernel: prepare -> FFT -> map -> post for every frame of one stream, one CTA per stream.
	// (included inside namespace sgz from sgz_spectrum.cu)
	//
	// FFT plan for N = 2 M, M = 16 * 16 * R3 (R3 = 16: N = 8192, R3 = 8: N = 4096), T = 16 R3 threads:
	// the complex input x[n] = (L w, R w)[n] is split into even and odd samples; thread u carries one
	// element of each (E, O) through three register-resident radix-16/16/R3 passes of two M-point
	// FFTs that share every twiddle; the exchanges between passes go through shared memory as float4
	// (E.re, E.im, O.re, O.im) with an XOR swizzle that keeps all three access patterns conflict free
	// and lets pass 2 run in place.  The last radix-2 (Z[k] = E[k] + W^k O[k], Z[k+M] = E[k] - W^k O[k])
	// is fused with the two-for-one split, which needs Z[k] and Z[N-k] = E[M-k] - W^(M-k) O[M-k].
	// (index algebra checked by tools/fft_plan_prototype.py)
	//
	// Latency plan: while the display-point phase of frame f runs, the lower half of the exchange buffer
	// is free, so the newest half of frame f+1's audio (the part that is not yet in L2) is brought in
	// with cp.async (LDGSTS, no registers); the older half is re-read from L2 at the top of the frame.
	// The barrier that protects the exchange buffer sits right before pass 1's first store, so warps
	// that finish their display points early already run the next frame's loads and first DFT16.
	// Per-point map tables live in shared memory; display points are processed two at a time so their
	// weight loads are in flight together.
	// ------------------------------------------------------------------------------------------
	struct FusedArgs
	{
		const float* audio; uint64_t nSamples; uint64_t streamStride; int layout;
		uint32_t hop, nFrames;
		uint32_t firstStream, numStreams; // the launch covers streams [firstStream, firstStream + numStreams)
		// k_fused8 only, frame-parallel mode: blockIdx.y selects a run of `framesPerChunk` frames (0 = all frames in one
		// CTA) and its own state slot - `states` is then [S][stateSlots][2][P], slot 0 holding the stream's real state and
		// the others zero; k_hold_fixup afterwards carries the peak hold from run to run
		uint32_t framesPerChunk, stateSlots;
		float2* states;  // [S][2][P]
		float2* results; // [S][F][2][P]
	};

	template<int R3> struct FusedGeom
	{
		static constexpr int T = 16 * R3;
		static constexpr int M = 16 * T;
		static constexpr int N = 2 * M;
		static constexpr int G = 16 / R3;
		static constexpr size_t xbufBytes = size_t(M) * 16;          // float4[M]; csfmag[N+1] aliases its upper half (+4 B)
		// small twiddle tables kept in shared memory: A[16][16] = W_256^(n2 k1), B[16][R3] = W_M^(n3 k1),
		// C[16][R3] = W_(16 R3)^(n3 k2)
		static constexpr size_t twOffset = xbufBytes + 16;
		static constexpr size_t twBytes = (256 + 2 * 16 * R3) * sizeof(float2);   // C is read as float2 and broadcast in registers
		static constexpr size_t stateOffset = twOffset + twBytes;
		// after the [2][P] float2 states: packed bin map (u32), slope (f32), Lanczos base of channel 2 (i32)
		__host__ __device__ static size_t mapOffset(uint32_t P) { return stateOffset + size_t(P) * 16; }
		static size_t smemBytes(uint32_t P) { return mapOffset(P) + size_t(P) * 12; }
	};

	template<int R3>
	__device__ __forceinline__ int physAddr(int k1, int j, int n3)
	{
		return k1 * (16 * R3) + j * R3 + (n3 ^ (k1 & 7));
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

	// packed bin map entry: bit 31 = max-of-bins, bits 30..16 = count, bits 15..0 = first
	__host__ __device__ __forceinline__ uint32_t packBinMap(int mode, int first, int count) { return (uint32_t(mode) << 31) | (uint32_t(count) << 16) | uint32_t(first); }

	// MATRIX = 0: (L w, R w) (Separate and the other two-channel inputs); 1: runtime channel matrix
	template<int R3, bool STEREO, int BINPOL, int MATRIX>
	__global__ void __launch_bounds__(16 * R3, (R3 == 16 ? 2 : 4))
	k_fused(DeviceTables t, FusedArgs a)
	{
		typedef FusedGeom<R3> Gm;
		constexpr int T = Gm::T, M = Gm::M, N = Gm::N;
		extern __shared__ __align__(16) unsigned char smemRaw[];
		float4* xbuf = reinterpret_cast<float4*>(smemRaw);
		float2* zm = reinterpret_cast<float2*>(smemRaw);                 // [M], lower half of xbuf
		float4* stage = reinterpret_cast<float4*>(smemRaw);               // [8][T] next frame's newest half (lower half of xbuf)
		float* csfmag = reinterpret_cast<float*>(smemRaw + size_t(M) * 8); // [N+1], upper half of xbuf (+4 B)
		float2* twA = reinterpret_cast<float2*>(smemRaw + Gm::twOffset);  // [16][16]  W_256^(n2 k1)
		float2* twB = twA + 256;                                          // [16][R3]  W_M^(n3 k1)
		float2* twC = twB + 16 * R3;                                      // [16][R3]  W_(16 R3)^(n3 k2)
		float2* st = reinterpret_cast<float2*>(smemRaw + Gm::stateOffset); // [2][P]

		const int u = threadIdx.x;
		const int stream = blockIdx.x + a.firstStream;
		const int P = int(t.P);
		const int W = int(t.W);
		const int conf = t.conf;
		const int linLimit = int(t.linLimit);
		uint32_t* mapPacked = reinterpret_cast<uint32_t*>(smemRaw + Gm::mapOffset(t.P));
		float* mapSlope = reinterpret_cast<float*>(mapPacked + P);
		int* mapBaseR = reinterpret_cast<int*>(mapSlope + P);

		// per-stream state and the per-view tables live in shared memory for the whole frame loop
		for (int i = u; i < 2 * P; i += T) st[i] = a.states[uint64_t(stream) * 2 * P + i];
		for (int i = u; i < P; i += T)
		{
			const int4 bm = t.binmap[i];
			mapPacked[i] = packBinMap(bm.x, bm.y, bm.z);
			mapSlope[i] = t.slope[i] * t.minFracRecip;  // slope * minFracRecip, folded once
			mapBaseR[i] = t.baseR[i];
		}
		for (int i = u; i < 256; i += T) twA[i] = t.tw1[i];
		for (int i = u; i < 16 * R3; i += T) { twB[i] = t.tw1[256 + i]; twC[i] = t.tw2[i]; }
		// W_N^(u + T g): the thread-constant part of the last radix-2 twiddle, two registers per g
		float2 wu[Gm::G];
#pragma unroll
		for (int g = 0; g < Gm::G; ++g) wu[g] = t.tw3[u + T * g];
		__syncthreads(); // tables are read (twA / twB in pass 1) before the first in-loop barrier

		const float* sbase = a.audio + uint64_t(stream) * a.streamStride;
		const bool vec4 = a.layout == SGZ_LAYOUT_INTERLEAVED && ((reinterpret_cast<uintptr_t>(sbase) & 15) == 0) && ((a.hop & 1) == 0) && W == N;
		bool staged = false; // the newest half of the current frame sits in `stage`

		for (uint32_t f = 0; f < a.nFrames; ++f)
		{
			// SoA register pairs: re[i] = (E.re, O.re), im[i] = (E.im, O.im)
			float2 re[16], im[16];
			const uint64_t start = uint64_t(f) * a.hop;

			// ---- pass 1: load, window + channel matrix, DFT16 over n1, twiddle W_M^(u k1) ----
			// raw = (L[2p], R[2p], L[2p+1], R[2p+1]) for sample pair p = n1 T + u
			// all loads of the frame are issued back to back before the first use (one L2 round trip, not 16)
			float2 wv[16];
#pragma unroll
			for (int n1 = 0; n1 < 16; ++n1) wv[n1] = ldg_f2(reinterpret_cast<const float2*>(t.window) + n1 * T + u);
			float4 rv[16];
			if (vec4)
			{
				const float4* src = reinterpret_cast<const float4*>(sbase + 2 * start) + u;
#pragma unroll
				for (int n1 = 0; n1 < 8; ++n1) rv[n1] = ldg_stream_f4(src + n1 * T);
				if (staged)
				{
					cpAsyncWaitAll(); // each thread reads back only what it staged itself
#pragma unroll
					for (int n1 = 8; n1 < 16; ++n1) rv[n1] = stage[(n1 - 8) * T + u];
				}
				else
				{
#pragma unroll
					for (int n1 = 8; n1 < 16; ++n1) rv[n1] = ldg_stream_f4(src + n1 * T);
				}
			}
			else
			{
#pragma unroll
				for (int n1 = 0; n1 < 16; ++n1)
				{
					const int n = 2 * (n1 * T + u);
					float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
					if (n < W) fetchLR(sbase, a.layout, a.nSamples, start + n, v.x, v.y);
					if (n + 1 < W) fetchLR(sbase, a.layout, a.nSamples, start + n + 1, v.z, v.w);
					rv[n1] = v;
				}
			}
#pragma unroll
			for (int n1 = 0; n1 < 16; ++n1)
			{
				const float4 v = rv[n1];
				const float2 w = wv[n1];
				if constexpr (MATRIX == 0)
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
			dft16p(re, im);
#pragma unroll
			for (int k1 = 1; k1 < 16; ++k1)
			{
				// W_M^(u k1) = W_256^(n2 k1) W_M^(n3 k1), u = n2 R3 + n3
				const float2 w = cmul(twA[k1 * 16 + u / R3], twB[k1 * R3 + u % R3]);
				mulTwiddleP(re[perm16(k1)], im[perm16(k1)], make_float4(w.x, w.x, w.y, w.y));
			}
			// everybody is done with the previous frame's csfmag and with the staged audio: the exchange
			// buffer may be overwritten (also orders the table fill before the first frame)
			__syncthreads();
#pragma unroll
			for (int k1 = 0; k1 < 16; ++k1)
			{
				const float2 r = re[perm16(k1)], i = im[perm16(k1)];
				xbuf[k1 * T + (u ^ (k1 & 7))] = make_float4(r.x, r.y, i.x, i.y); // physAddr(k1, n2, n3), u = n2 R3 + n3
			}
			__syncthreads();

			// ---- pass 2 (in place): DFT16 over n2, twiddle W_(16 R3)^(n3 k2) ----
			{
				const int k1 = u / R3, n3 = u % R3;
#pragma unroll
				for (int n2 = 0; n2 < 16; ++n2)
				{
					const float4 v = xbuf[physAddr<R3>(k1, n2, n3)];
					re[n2] = make_float2(v.x, v.y); im[n2] = make_float2(v.z, v.w);
				}
				dft16p(re, im);
#pragma unroll
				for (int k2 = 0; k2 < 16; ++k2)
				{
					float2 r = re[perm16(k2)], i = im[perm16(k2)];
					if (k2 > 0)
					{
						const float2 w = twC[k2 * R3 + n3];
						mulTwiddleP(r, i, make_float4(w.x, w.x, w.y, w.y));
					}
					xbuf[physAddr<R3>(k1, k2, n3)] = make_float4(r.x, r.y, i.x, i.y);
				}
			}
			__syncthreads();

			// ---- pass 3: DFT_R3 over n3; thread owns bins k = u + T g + 256 k3 ----
#pragma unroll
			for (int g = 0; g < Gm::G; ++g)
			{
				const int q = u + T * g;
				const int k1 = q & 15, k2 = q >> 4;
#pragma unroll
				for (int n3 = 0; n3 < R3; ++n3)
				{
					const float4 v = xbuf[physAddr<R3>(k1, k2, n3)];
					re[g * R3 + n3] = make_float2(v.x, v.y); im[g * R3 + n3] = make_float2(v.z, v.w);
				}
			}
			if constexpr (R3 == 16) dft16p(re, im);
			else { dft8p<0>(re, im); dft8p<8>(re, im); }

			// ---- last radix-2: Zp[k] = E + W_N^k O (= Z[k]), Zm[k] = E - W_N^k O (= Z[k + M]) ----
			// afterwards re[slot] = (Zp.re, Zm.re), im[slot] = (Zp.im, Zm.im), slot = g R3 + perm(k3)
#pragma unroll
			for (int g = 0; g < Gm::G; ++g)
			{
#pragma unroll
				for (int k3 = 0; k3 < R3; ++k3)
				{
					const int slot = g * R3 + (R3 == 16 ? perm16(k3) : perm8(k3));
					// W_N^k, k = u + T g + 256 k3: W_N^256 = W_(2 R3)
					const float2 o = make_float2(re[slot].y, im[slot].y), e = make_float2(re[slot].x, im[slot].x);
					const float2 wo = cmul(cmul(o, w32(R3 == 16 ? k3 : 2 * k3)), wu[g]);
					re[slot] = make_float2(e.x + wo.x, e.x - wo.x);
					im[slot] = make_float2(e.y + wo.y, e.y - wo.y);
				}
			}
			__syncthreads(); // every pass-3 read of xbuf is done before zm / csfmag overwrite it

			if constexpr (STEREO)
			{
#pragma unroll
				for (int g = 0; g < Gm::G; ++g)
#pragma unroll
					for (int k3 = 0; k3 < R3; ++k3)
					{
						const int slot = g * R3 + (R3 == 16 ? perm16(k3) : perm8(k3));
						zm[u + T * g + 256 * k3] = make_float2(re[slot].y, im[slot].y);
					}
				__syncthreads();
				// two-for-one split + |.| (or |.|^2 beyond linLimit) into the reference's csf layout:
				// channel 1 bin j at csfmag[j], channel 2 bin j at csfmag[N - j]
#pragma unroll
				for (int g = 0; g < Gm::G; ++g)
#pragma unroll
					for (int k3 = 0; k3 < R3; ++k3)
					{
						const int slot = g * R3 + (R3 == 16 ? perm16(k3) : perm8(k3));
						const int j = u + T * g + 256 * k3;
						const float2 zj = make_float2(re[slot].x, im[slot].x);
						if (j == 0)
						{
							// :861-864: csf[N] = Im Z[0] / 2, csf[0] = Re Z[0] / 2, csf[N/2] *= .5 (packed Nyquists)
							const float2 ny = make_float2(re[slot].y, im[slot].y); // Z[M] = Zm[0]
							const float h2 = 0.25f * cnorm(ny);
							csfmag[0] = zj.x * 0.5f;   // linLimit >= 1: entries 0 and N are always magnitudes
							csfmag[N] = zj.y * 0.5f;
							csfmag[M] = (M < linLimit) ? approxSqrt(h2) : h2;
							continue;
						}
						const float2 zn = zm[M - j]; // Z[N - j]
						float ax = (zj.x + zn.x) * 0.5f, ay = (zj.y - zn.y) * 0.5f;
						const float dx = (zj.x - zn.x) * 0.5f, dy = (zj.y + zn.y) * 0.5f;
						if (j == M - 1) { ax *= 0.5f; ay *= 0.5f; } // csf[(N >> 1) - 1] *= 0.5, channel 1 only
						float a2 = ax * ax + ay * ay, b2 = dx * dx + dy * dy;
						if (j < linLimit) { a2 = approxSqrt(a2); b2 = approxSqrt(b2); }
						csfmag[j] = a2;
						csfmag[N - j] = b2;
					}
			}
			else
			{
				// mono: csf[i] = |Z[i]| for i < M (csf[0] halved), csf[M] = Z[M] / 2 (kept as |.|)
#pragma unroll
				for (int g = 0; g < Gm::G; ++g)
#pragma unroll
					for (int k3 = 0; k3 < R3; ++k3)
					{
						const int slot = g * R3 + (R3 == 16 ? perm16(k3) : perm8(k3));
						const int j = u + T * g + 256 * k3;
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
			__syncthreads();

			// the lower half of the exchange buffer (zm) is dead until the next frame's pass 1: bring in
			// the newest half of the next frame's audio behind the display-point phase
			staged = false;
			if (vec4 && f + 1 < a.nFrames)
			{
				const float4* src = reinterpret_cast<const float4*>(sbase + 2 * (start + a.hop)) + u;
#pragma unroll
				for (int n1 = 8; n1 < 16; ++n1) cpAsync16(stage + (n1 - 8) * T + u, src + n1 * T);
				cpAsyncCommit();
				staged = true;
			}

			// ---- display-point mapping + log scale + peak hold, two points per thread at a time ----
			float2* resOut = a.results + ((uint64_t(stream) * a.nFrames + f) * 2) * P;
			const float poles[2] = { t.pole0, t.pole1 };
			const int safeFrom = int(t.safeFrom); // interpolated points x >= safeFrom have every tap in range
			for (int x0 = 0; x0 < P; x0 += 2 * T)
			{
				uint32_t pkd[2]; float4 wl[2][3], wr[2][3]; int br[2];
#pragma unroll
				for (int i = 0; i < 2; ++i)
				{
					const int x = x0 + i * T + u;
					pkd[i] = x < P ? mapPacked[x] : 0xFFFFFFFFu;
					if constexpr (BINPOL == SGZ_BIN_LANCZOS)
					{
						if (!(pkd[i] >> 31))
						{
							// weights are stored 12 floats per point (10 taps + 2 zeros) so three float4 fetch them
							const float4* wl4 = reinterpret_cast<const float4*>(t.wL + size_t(x) * kLanczosStride);
							wl[i][0] = __ldg(wl4); wl[i][1] = __ldg(wl4 + 1); wl[i][2] = __ldg(wl4 + 2);
							if constexpr (STEREO)
							{
								const float4* wr4 = reinterpret_cast<const float4*>(t.wR + size_t(x) * kLanczosStride);
								wr[i][0] = __ldg(wr4); wr[i][1] = __ldg(wr4 + 1); wr[i][2] = __ldg(wr4 + 2);
								br[i] = mapBaseR[x];
							}
						}
					}
				}
#pragma unroll
				for (int i = 0; i < 2; ++i)
				{
					const int x = x0 + i * T + u;
					if (x >= P) continue;
					const int first = int(pkd[i] & 0xFFFFu), count = int((pkd[i] >> 16) & 0x7FFFu);
					float vL, vR = 0.f;
					if (!(pkd[i] >> 31))
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
							}
							if constexpr (STEREO)
							{
								const float pos = __ldg(t.posR + x); const float fl = floorf(pos); const int k = int(fl); const float fr = pos - fl;
								vR = csfTap(csfmag, k, N) * (1.f - fr) + csfTap(csfmag, k + 1, N) * fr;
							}
						}
						else
						{
							const int bl = first - (kLanczosSize - 1); // first = floor(pos)
							const float w[10] = { wl[i][0].x, wl[i][0].y, wl[i][0].z, wl[i][0].w, wl[i][1].x, wl[i][1].y, wl[i][1].z, wl[i][1].w, wl[i][2].x, wl[i][2].y };
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
							if constexpr (STEREO)
							{
								const float v[10] = { wr[i][0].x, wr[i][0].y, wr[i][0].z, wr[i][0].w, wr[i][1].x, wr[i][1].y, wr[i][1].z, wr[i][1].w, wr[i][2].x, wr[i][2].y };
								if (x >= safeFrom)
								{
#pragma unroll
									for (int k = 0; k < kLanczosTaps; ++k) vR += csfmag[br[i] + k] * v[k];
								}
								else
								{
#pragma unroll
									for (int k = 0; k < kLanczosTaps; ++k) vR += csfTap(csfmag, br[i] + k, N) * v[k];
								}
							}
						}
					}
					else
					{
						// max-of-bins (:944-984).  The display value is the magnitude at the arg-max, so the
						// running maximum of the (squared) entries is all that is needed; when nothing exceeds
						// zero the reference falls back to csf[bin] for BOTH channels (maxRBin = maxLBin = bin).
						float bestL = 0.f, bestR = 0.f;
						if (first >= linLimit)
						{
							for (int c = 0; c < count; ++c)
							{
								bestL = fmaxf(bestL, csfmag[first + c]);
								if constexpr (STEREO) bestR = fmaxf(bestR, csfmag[N - first - c]);
							}
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
						const int fb = first + count - 1; // == the reference's `bin`
						const float fv = csfmag[fb];
						const float fbMag = (fb < linLimit || fb > N - linLimit) ? fv : approxSqrt(fv);
						vL = bestL > 0.f ? approxSqrt(bestL) : fbMag;
						if constexpr (STEREO) vR = bestR > 0.f ? approxSqrt(bestR) : fbMag;
					}
					const float lmag = fabsf(t.invSize * vL);
					const float rmag = fabsf(t.invSize * vR);
					const float sm = mapSlope[x];
#pragma unroll
					for (int k = 0; k < 2; ++k)
					{
						float2 sv = st[k * P + x];
						sv.x *= poles[k];
						if (lmag > sv.x) sv.x = lmag;
						float2 res;
						const float dl = sm * sv.x;
						res.x = dl > 0.f ? __log2f(dl) * t.deltaYLn2 : t.lowerClip;
						if constexpr (STEREO)
						{
							sv.y *= poles[k];
							if (rmag > sv.y) sv.y = rmag;
							const float dr = sm * sv.y;
							res.y = dr > 0.f ? __log2f(dr) * t.deltaYLn2 : t.lowerClip;
						}
						else res.y = 0.f;
						st[k * P + x] = sv;
						stg_stream_f2(resOut + k * P + x, res);
					}
				}
			}
			// no barrier here: the next frame's pass 1 synchronises before it touches the exchange buffer
		}
		__syncthreads();
		for (int i = u; i < 2 * P; i += T) a.states[uint64_t(stream) * 2 * P + i] = st[i];
	}
