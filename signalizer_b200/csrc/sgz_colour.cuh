	// ------------------------------------------------------------------------------------------
	// Oscilloscope colouring and envelopes (included inside namespace sgz from sgz_scope.cu).
	//
	// k_scope_colour: the per-sample loop of StreamState::audioProcessing (OscilloscopeDSP.inl:505-651).
	// Everything here is a chain of recursive filters, so time cannot be split without changing the
	// rounding; the recursions run on two lanes of a warp (left / right channel of one stereo pair: that
	// channel's crossover network and colour smoothing, the three band outputs swapped with one shuffle each
	// for the mid / side states) 32 samples at a time into shared memory; accumulateColour, which is
	// feed-forward, is then done by the whole warp - lane j colours sample j - and each of the four planes
	// gets one 128-byte row per chunk.  Every float operation is an explicitly rounded one (no fused
	// multiply-add) so that the bytes equal the CPU restatement's.  Phase 1 is latency-bound by construction
	// (one dependent section chain per sample): the GPU is only filled from several thousand stereo pairs up.
	// ------------------------------------------------------------------------------------------
	struct ColourSection { float b0, b1, b2, a1, a2; };
	struct ColourArgs
	{
		const float* audio; int layout; uint64_t n; uint64_t stride; uint32_t nStreams;
		ColourSection lp[2], hp[2];
		float pole, envelopeCoeff, blend;
		float colours[3][3];
		float keyL[4], keyR[4];
		int channelMode, envelopeMode;
		sgz_scope_colour_state* states; // [S][2] device copy
		uint8_t* rgba;                  // [S][4][n][4]
	};

	__device__ __forceinline__ float mulr(float a, float b) { return __fmul_rn(a, b); }
	__device__ __forceinline__ float addr(float a, float b) { return __fadd_rn(a, b); }
	__device__ __forceinline__ float subr(float a, float b) { return __fsub_rn(a, b); }

	__device__ __forceinline__ float sectionStep(const ColourSection& s, float& z0, float& z1, float x)
	{
		const float y = addr(mulr(s.b0, x), z0);
		z0 = addr(subr(mulr(s.b1, x), mulr(s.a1, y)), z1);
		z1 = subr(mulr(s.b2, x), mulr(s.a2, y));
		return y;
	}

	__device__ __forceinline__ void smoothSquares(const float (&bands)[3], float (&st)[3], float pole)
	{
#pragma unroll
		for (int i = 0; i < 3; ++i)
		{
			const float in = mulr(bands[i], bands[i]);
			st[i] = addr(in, mulr(pole, subr(st[i], in)));
		}
	}

	__device__ __forceinline__ uint32_t colourOf(const float (&st)[3], const float (&col)[3][3], const float* key, float blend)
	{
		float rgb[3];
#pragma unroll
		for (int c = 0; c < 3; ++c)
		{
			float acc = 0.f;
#pragma unroll
			for (int i = 0; i < 3; ++i) acc = addr(acc, mulr(st[i], col[i][c]));
			rgb[c] = acc;
		}
		const float invMax = __fdiv_rn(255.0f, fmaxf(rgb[0], fmaxf(rgb[2], rgb[1])));
		uint32_t out = 0;
#pragma unroll
		for (int c = 0; c < 4; ++c)
		{
			// float -> byte truncates; NaN (silence: 0 * inf) becomes 0
			const float px = c < 3 ? float(__float2uint_rz(mulr(rgb[c], invMax)) & 0xffu) : 255.0f;
			const uint32_t b = __float2uint_rz(addr(px, mulr(subr(key[c], px), blend))) & 0xffu;
			out |= b << (8 * c);
		}
		return out;
	}

	constexpr int kColourChunk = 32;        // samples filtered by lanes 0 / 1 before the whole warp colours them
	constexpr int kColourRow = 13;          // 12 smoothed band energies per sample + 1 pad word (conflict-free rows)
	constexpr int kColourWarps = 4;
	constexpr uint32_t kColourPackedFrom = 2048; // stereo pairs from which k_scope_colour_packed takes over

	__global__ void __launch_bounds__(32 * kColourWarps) k_scope_colour(ColourArgs a)
	{
		__shared__ float smoothed[kColourWarps][kColourChunk][kColourRow]; // [left 3 | right 3 | mid 3 | side 3]
		const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
		const uint32_t s = blockIdx.x * kColourWarps + warp; // one warp per stereo pair
		if (s >= a.nStreams) return;
		const int ch = lane & 1;
		const float* src = a.audio + uint64_t(s) * a.stride;
		sgz_scope_colour_state* st = a.states + uint64_t(s) * 2 + ch;
		float z[24], sm[3], aux[3], env = 0.f;
		if (lane < 2)
		{
#pragma unroll
			for (int i = 0; i < 24; ++i) z[i] = st->network[i];
#pragma unroll
			for (int i = 0; i < 3; ++i) { sm[i] = st->smooth_filters[i]; aux[i] = st->aux_smooth_filter[i]; }
			// the single-trace modes keep one envelope (channel 0's) and copy it to the other channel (:521-556)
			env = (a.channelMode <= SGZ_OSC_SIDE) ? a.states[uint64_t(s) * 2].envelope : st->envelope;
		}
		uint32_t* planes = reinterpret_cast<uint32_t*>(a.rgba) + uint64_t(s) * 4 * a.n;
		for (uint64_t i0 = 0; i0 < a.n; i0 += kColourChunk)
		{
			const int count = int(min(uint64_t(kColourChunk), a.n - i0));
			if (lane < 2)
			{
				// phase 1: the recursions, sample by sample, on the two channel lanes
				for (int j = 0; j < count; ++j)
				{
					float l, r;
					loadLR(src, a.layout, a.n, i0 + j, l, r);
					if (a.envelopeMode != SGZ_ENVELOPE_NONE)
					{
						float sample;
						switch (a.channelMode)
						{
						case SGZ_OSC_LEFT: sample = mulr(l, l); break;
						case SGZ_OSC_RIGHT: sample = mulr(r, r); break;
						case SGZ_OSC_MID: { const float m = mulr(0.5f, addr(l, r)); sample = mulr(m, m); break; }
						case SGZ_OSC_SIDE: { const float d = mulr(0.5f, subr(l, r)); sample = mulr(d, d); break; }
						case SGZ_OSC_SEPARATE: { const float x = ch == 0 ? l : r; sample = mulr(x, x); break; }
						default: { const float x = ch == 0 ? addr(l, r) : subr(l, r); sample = mulr(0.5f, mulr(x, x)); break; }
						}
						env = addr(sample, mulr(a.envelopeCoeff, subr(env, sample)));
					}
					const float x = ch == 0 ? l : r;
					float bands[3], ms[3];
					{
						const float low = sectionStep(a.lp[0], z[2], z[3], sectionStep(a.lp[0], z[0], z[1], x));
						const float rest = sectionStep(a.hp[0], z[6], z[7], sectionStep(a.hp[0], z[4], z[5], x));
						const float apl = sectionStep(a.lp[1], z[10], z[11], sectionStep(a.lp[1], z[8], z[9], low));
						const float aph = sectionStep(a.hp[1], z[14], z[15], sectionStep(a.hp[1], z[12], z[13], low));
						bands[0] = addr(apl, aph);
						bands[1] = sectionStep(a.lp[1], z[18], z[19], sectionStep(a.lp[1], z[16], z[17], rest));
						bands[2] = sectionStep(a.hp[1], z[22], z[23], sectionStep(a.hp[1], z[20], z[21], rest));
					}
					smoothSquares(bands, sm, a.pole);
					// lane 0: mid = left + right; lane 1: side = left - right
#pragma unroll
					for (int k = 0; k < 3; ++k)
					{
						const float other = __shfl_xor_sync(0x3u, bands[k], 1);
						ms[k] = ch == 0 ? addr(bands[k], other) : subr(other, bands[k]);
					}
					smoothSquares(ms, aux, a.pole);
					float* row = smoothed[warp][j];
#pragma unroll
					for (int k = 0; k < 3; ++k) { row[3 * ch + k] = sm[k]; row[6 + 3 * ch + k] = aux[k]; }
				}
			}
			__syncwarp();
			// phase 2: accumulateColour is feed-forward: lane j colours sample j of the chunk, 128-byte rows per plane
			if (lane < count)
			{
				const float* row = smoothed[warp][lane];
#pragma unroll
				for (int p = 0; p < 4; ++p)
				{
					const float e[3] = { row[3 * p], row[3 * p + 1], row[3 * p + 2] };
					planes[uint64_t(p) * a.n + i0 + lane] = colourOf(e, a.colours, (p & 1) ? a.keyR : a.keyL, a.blend);
				}
			}
			__syncwarp();
		}
		if (lane >= 2) return;
#pragma unroll
		for (int i = 0; i < 24; ++i) st->network[i] = z[i];
#pragma unroll
		for (int i = 0; i < 3; ++i) { st->smooth_filters[i] = sm[i]; st->aux_smooth_filter[i] = aux[i]; }
		if (a.envelopeMode == SGZ_ENVELOPE_RMS) st->envelope = env;
	}

	// Many-stream variant: with thousands of stereo pairs the recursions themselves fill the machine, so every lane
	// pair (2j, 2j+1) of a warp owns one stereo pair and also colours its own two planes (left + mid / right + side),
	// four pixels per 16-byte store.  Same arithmetic, statement for statement, as k_scope_colour.
	__global__ void __launch_bounds__(64) k_scope_colour_packed(ColourArgs a)
	{
		const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t s = t >> 1;
		const int ch = t & 1;
		const bool live = s < a.nStreams;
		const uint32_t ss = live ? s : a.nStreams - 1; // idle lanes shadow the last pair so the shuffles stay uniform
		const float* src = a.audio + uint64_t(ss) * a.stride;
		sgz_scope_colour_state* st = a.states + uint64_t(ss) * 2 + ch;
		float z[24], sm[3], aux[3];
#pragma unroll
		for (int i = 0; i < 24; ++i) z[i] = st->network[i];
#pragma unroll
		for (int i = 0; i < 3; ++i) { sm[i] = st->smooth_filters[i]; aux[i] = st->aux_smooth_filter[i]; }
		float env = (a.channelMode <= SGZ_OSC_SIDE) ? a.states[uint64_t(ss) * 2].envelope : st->envelope;
		const float* key = ch == 0 ? a.keyL : a.keyR;
		uint32_t* planeOwn = reinterpret_cast<uint32_t*>(a.rgba) + (uint64_t(ss) * 4 + ch) * a.n;     // left | right
		uint32_t* planeAux = reinterpret_cast<uint32_t*>(a.rgba) + (uint64_t(ss) * 4 + 2 + ch) * a.n; // mid | side
		for (uint64_t i0 = 0; i0 < a.n; i0 += 4)
		{
			uint32_t own4[4], aux4[4];
#pragma unroll
			for (int j = 0; j < 4; ++j)
			{
				const uint64_t i = i0 + j;
				if (i >= a.n) break; // uniform: every lane runs the same n
				float l, r;
				loadLR(src, a.layout, a.n, i, l, r);
				if (a.envelopeMode != SGZ_ENVELOPE_NONE)
				{
					float sample;
					switch (a.channelMode)
					{
					case SGZ_OSC_LEFT: sample = mulr(l, l); break;
					case SGZ_OSC_RIGHT: sample = mulr(r, r); break;
					case SGZ_OSC_MID: { const float m = mulr(0.5f, addr(l, r)); sample = mulr(m, m); break; }
					case SGZ_OSC_SIDE: { const float d = mulr(0.5f, subr(l, r)); sample = mulr(d, d); break; }
					case SGZ_OSC_SEPARATE: { const float x = ch == 0 ? l : r; sample = mulr(x, x); break; }
					default: { const float x = ch == 0 ? addr(l, r) : subr(l, r); sample = mulr(0.5f, mulr(x, x)); break; }
					}
					env = addr(sample, mulr(a.envelopeCoeff, subr(env, sample)));
				}
				const float x = ch == 0 ? l : r;
				float bands[3], ms[3];
				{
					const float low = sectionStep(a.lp[0], z[2], z[3], sectionStep(a.lp[0], z[0], z[1], x));
					const float rest = sectionStep(a.hp[0], z[6], z[7], sectionStep(a.hp[0], z[4], z[5], x));
					const float apl = sectionStep(a.lp[1], z[10], z[11], sectionStep(a.lp[1], z[8], z[9], low));
					const float aph = sectionStep(a.hp[1], z[14], z[15], sectionStep(a.hp[1], z[12], z[13], low));
					bands[0] = addr(apl, aph);
					bands[1] = sectionStep(a.lp[1], z[18], z[19], sectionStep(a.lp[1], z[16], z[17], rest));
					bands[2] = sectionStep(a.hp[1], z[22], z[23], sectionStep(a.hp[1], z[20], z[21], rest));
				}
				smoothSquares(bands, sm, a.pole);
#pragma unroll
				for (int k = 0; k < 3; ++k)
				{
					const float other = __shfl_xor_sync(0xffffffffu, bands[k], 1);
					ms[k] = ch == 0 ? addr(bands[k], other) : subr(other, bands[k]);
				}
				smoothSquares(ms, aux, a.pole);
				own4[j] = colourOf(sm, a.colours, key, a.blend);
				aux4[j] = colourOf(aux, a.colours, key, a.blend);
			}
			if (live)
			{
				if (i0 + 4 <= a.n && ((reinterpret_cast<uintptr_t>(planeOwn + i0) | reinterpret_cast<uintptr_t>(planeAux + i0)) & 15) == 0)
				{
					*reinterpret_cast<uint4*>(planeOwn + i0) = make_uint4(own4[0], own4[1], own4[2], own4[3]);
					*reinterpret_cast<uint4*>(planeAux + i0) = make_uint4(aux4[0], aux4[1], aux4[2], aux4[3]);
				}
				else
				{
#pragma unroll
					for (int j = 0; j < 4; ++j) if (i0 + j < a.n) { planeOwn[i0 + j] = own4[j]; planeAux[i0 + j] = aux4[j]; }
				}
			}
		}
		if (!live) return;
#pragma unroll
		for (int i = 0; i < 24; ++i) st->network[i] = z[i];
#pragma unroll
		for (int i = 0; i < 3; ++i) { st->smooth_filters[i] = sm[i]; st->aux_smooth_filter[i] = aux[i]; }
		if (a.envelopeMode == SGZ_ENVELOPE_RMS) st->envelope = env;
	}

	// runPeakFilter's window maxima (OscilloscopeDSP.inl:746-867): one CTA per stream; out[s] = (max |a|, max |b|)
	// with (a, b) the two traces of the channel mode (single-trace modes: b = a)
	__global__ void __launch_bounds__(256) k_scope_peak_max(const float* __restrict__ audio, int layout, uint64_t n, uint64_t stride, uint64_t stop, int channelMode,
		float2* __restrict__ out)
	{
		const float* src = audio + uint64_t(blockIdx.x) * stride;
		float ma = 0.f, mb = 0.f;
		for (uint64_t i = threadIdx.x; i < stop; i += blockDim.x)
		{
			float l, r;
			loadLR(src, layout, n, i, l, r);
			float va, vb;
			switch (channelMode)
			{
			case SGZ_OSC_LEFT: va = vb = l; break;
			case SGZ_OSC_RIGHT: va = vb = r; break;
			case SGZ_OSC_MID: va = vb = mulr(addr(l, r), 0.5f); break;
			case SGZ_OSC_SIDE: va = vb = mulr(subr(l, r), 0.5f); break;
			case SGZ_OSC_SEPARATE: va = l; vb = r; break;
			default: va = mulr(addr(l, r), 0.5f); vb = mulr(subr(l, r), 0.5f); break;
			}
			ma = fmaxf(ma, fabsf(va)); mb = fmaxf(mb, fabsf(vb));
		}
		__shared__ float2 part[8];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) { ma = fmaxf(ma, __shfl_xor_sync(0xffffffffu, ma, o)); mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o)); }
		if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = make_float2(ma, mb);
		__syncthreads();
		if (threadIdx.x == 0)
		{
			for (int w = 1; w < 8; ++w) { ma = fmaxf(ma, part[w].x); mb = fmaxf(mb, part[w].y); }
			out[blockIdx.x] = make_float2(ma, mb);
		}
	}
