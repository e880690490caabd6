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
	// Latency plan: the audio of frame f+1 is loaded into the (then dead) FFT registers right before
	// the display-point phase of frame f, so DRAM/L2 latency hides behind that phase; display points
	// are processed four at a time per thread so their table loads are in flight together.
	// ------------------------------------------------------------------------------------------
	struct FusedArgs
	{
		const float* audio; uint64_t nSamples; uint64_t streamStride; int layout;
		uint32_t hop, nFrames;
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
		static size_t smemBytes(uint32_t P) { return stateOffset + size_t(P) * 16; }
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
		float* csfmag = reinterpret_cast<float*>(smemRaw + size_t(M) * 8); // [N+1], upper half of xbuf (+4 B)
		float2* twA = reinterpret_cast<float2*>(smemRaw + Gm::twOffset);  // [16][16]  W_256^(n2 k1)
		float2* twB = twA + 256;                                          // [16][R3]  W_M^(n3 k1)
		float2* twC = twB + 16 * R3;                                      // [16][R3]  W_(16 R3)^(n3 k2)
		float2* st = reinterpret_cast<float2*>(smemRaw + Gm::stateOffset); // [2][P]

		const int u = threadIdx.x;
		const int stream = blockIdx.x;
		const int P = int(t.P);
		const int W = int(t.W);
		const int conf = t.conf;
		const int linLimit = int(t.linLimit);

		// peak-hold state of this stream lives in shared memory for the whole frame loop
		for (int i = u; i < 2 * P; i += T) st[i] = a.states[uint64_t(stream) * 2 * P + i];
		for (int i = u; i < 256; i += T) twA[i] = t.tw1[i];
		for (int i = u; i < 16 * R3; i += T) { twB[i] = t.tw1[256 + i]; twC[i] = t.tw2[i]; }
		// W_N^(u + T g): the thread-constant part of the last radix-2 twiddle, two registers per g
		float2 wu[Gm::G];
#pragma unroll
		for (int g = 0; g < Gm::G; ++g) wu[g] = t.tw3[u + T * g];
		__syncthreads();

		const float* sbase = a.audio + uint64_t(stream) * a.streamStride;
		const bool vec4 = a.layout == SGZ_LAYOUT_INTERLEAVED && ((reinterpret_cast<uintptr_t>(sbase) & 15) == 0) && ((a.hop & 1) == 0);

		// raw[n1] = (L[2p], R[2p], L[2p+1], R[2p+1]) for sample pair p = n1 T + u of the frame
		float4 raw[16];
		auto loadFrame = [&](uint32_t f)
		{
			const uint64_t start = uint64_t(f) * a.hop;
			if (vec4 && W == N)
			{
				const float4* src = reinterpret_cast<const float4*>(sbase + 2 * start) + u;
#pragma unroll
				for (int n1 = 0; n1 < 16; ++n1) raw[n1] = ldg_stream_f4(src + n1 * T);
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
					raw[n1] = v;
				}
			}
		};
		loadFrame(0);

		for (uint32_t f = 0; f < a.nFrames; ++f)
		{
			// SoA register pairs: re[i] = (E.re, O.re), im[i] = (E.im, O.im)
			float2 re[16], im[16];

			// ---- pass 1: window + channel matrix, DFT16 over n1, twiddle W_M^(u k1) ----
#pragma unroll
			for (int n1 = 0; n1 < 16; ++n1)
			{
				const float2 w = ldg_f2(reinterpret_cast<const float2*>(t.window) + n1 * T + u);
				const float4 v = raw[n1];
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
			for (int k1 = 0; k1 < 16; ++k1)
			{
				float2 r = re[perm16(k1)], i = im[perm16(k1)];
				if (k1 > 0)
				{
					// W_M^(u k1) = W_256^(n2 k1) W_M^(n3 k1), u = n2 R3 + n3
					const float2 w = cmul(twA[k1 * 16 + u / R3], twB[k1 * R3 + u % R3]);
					mulTwiddleP(r, i, make_float4(w.x, w.x, w.y, w.y));
				}
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
							const float h = 0.5f * cabs(ny);
							csfmag[0] = zj.x * 0.5f;   // linLimit >= 1: entries 0 and N are always magnitudes
							csfmag[N] = zj.y * 0.5f;
							csfmag[M] = (M < linLimit) ? h : h * h;
							continue;
						}
						const float2 zn = zm[M - j]; // Z[N - j]
						float ax = (zj.x + zn.x) * 0.5f, ay = (zj.y - zn.y) * 0.5f;
						const float dx = (zj.x - zn.x) * 0.5f, dy = (zj.y + zn.y) * 0.5f;
						if (j == M - 1) { ax *= 0.5f; ay *= 0.5f; } // csf[(N >> 1) - 1] *= 0.5, channel 1 only
						float a2 = ax * ax + ay * ay, b2 = dx * dx + dy * dy;
						if (j < linLimit) { a2 = sqrtf(a2); b2 = sqrtf(b2); }
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
							csfmag[M] = (M < linLimit) ? sqrtf(h2) : h2;
						}
						const float m2 = cnorm(zj);
						csfmag[j] = (j < linLimit) ? sqrtf(m2) : m2;
					}
			}

			// the FFT registers are dead from here to the end of the frame: fetch the next frame now
			if (f + 1 < a.nFrames) loadFrame(f + 1);
			__syncthreads();

			// ---- display-point mapping + log scale + peak hold, four points per thread at a time ----
			float2* resOut = a.results + ((uint64_t(stream) * a.nFrames + f) * 2) * P;
			const float poles[2] = { t.pole0, t.pole1 };
			for (int x0 = 0; x0 < P; x0 += 4 * T)
			{
				int4 bm[4]; float slope[4];
#pragma unroll
				for (int i = 0; i < 4; ++i)
				{
					const int x = x0 + i * T + u;
					const int xc = x < P ? x : P - 1;
					bm[i] = __ldg(t.binmap + xc);
					slope[i] = __ldg(t.slope + xc);
				}
#pragma unroll
				for (int i = 0; i < 4; ++i)
				{
					const int x = x0 + i * T + u;
					if (x >= P) continue;
					float vL, vR = 0.f;
					if (bm[i].x == 0)
					{
						if constexpr (BINPOL == SGZ_BIN_NONE)
						{
							vL = csfmag[bm[i].y];
							if constexpr (STEREO) vR = csfmag[N - bm[i].y];
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
							// weights are stored 12 floats per point (10 taps + 2 zeros) so three float4 fetch them
							const int bl = __ldg(t.baseL + x);
							const float4* wl4 = reinterpret_cast<const float4*>(t.wL + size_t(x) * kLanczosStride);
							const float4 w0 = __ldg(wl4), w1 = __ldg(wl4 + 1), w2 = __ldg(wl4 + 2);
							float4 r0, r1, r2; int br = 0;
							if constexpr (STEREO)
							{
								br = __ldg(t.baseR + x);
								const float4* wr4 = reinterpret_cast<const float4*>(t.wR + size_t(x) * kLanczosStride);
								r0 = __ldg(wr4); r1 = __ldg(wr4 + 1); r2 = __ldg(wr4 + 2);
							}
							const float wl[10] = { w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y };
							vL = 0.f;
#pragma unroll
							for (int k = 0; k < kLanczosTaps; ++k) vL += csfTap(csfmag, bl + k, N) * wl[k];
							if constexpr (STEREO)
							{
								const float wr[10] = { r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y };
#pragma unroll
								for (int k = 0; k < kLanczosTaps; ++k) vR += csfTap(csfmag, br + k, N) * wr[k];
							}
						}
					}
					else
					{
						// max-of-bins (:944-984): entries beyond linLimit are already squared
						int maxL = bm[i].w, maxR = bm[i].w; float bestL = 0.f, bestR = 0.f;
						for (int c = 0; c < bm[i].z; ++c)
						{
							const int off = bm[i].y + c;
							const float v = csfmag[off];
							const float mL = (off < linLimit) ? v * v : v;
							if (mL > bestL) { bestL = mL; maxL = off; }
							if constexpr (STEREO)
							{
								const float vr = csfmag[N - off];
								const float mR = (off < linLimit) ? vr * vr : vr;
								if (mR > bestR) { bestR = mR; maxR = N - off; }
							}
						}
						{
							const float v = csfmag[maxL];
							vL = (maxL < linLimit || maxL > N - linLimit) ? v : sqrtf(v);
						}
						if constexpr (STEREO)
						{
							const float v = csfmag[maxR];
							vR = (maxR < linLimit || maxR > N - linLimit) ? v : sqrtf(v);
						}
					}
					const float lmag = fabsf(t.invSize * vL);
					const float rmag = fabsf(t.invSize * vR);
#pragma unroll
					for (int k = 0; k < 2; ++k)
					{
						float2 sv = st[k * P + x];
						sv.x *= poles[k];
						if (lmag > sv.x) sv.x = lmag;
						float2 res;
						res.x = fastLogScale(slope[i], sv.x, t.minFracRecip, t.deltaYLn2, t.lowerClip);
						if constexpr (STEREO)
						{
							sv.y *= poles[k];
							if (rmag > sv.y) sv.y = rmag;
							res.y = fastLogScale(slope[i], sv.y, t.minFracRecip, t.deltaYLn2, t.lowerClip);
						}
						else res.y = 0.f;
						st[k * P + x] = sv;
						stg_stream_f2(resOut + k * P + x, res);
					}
				}
			}
			__syncthreads(); // csfmag / zm are overwritten by the next frame's pass 1
		}

		for (int i = u; i < 2 * P; i += T) a.states[uint64_t(stream) * 2 * P + i] = st[i];
	}
