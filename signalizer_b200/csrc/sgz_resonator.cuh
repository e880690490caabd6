	// ------------------------------------------------------------------------------------------
	// Resonator bank (RSNT algorithm) — included inside namespace sgz from sgz_spectrum.cu.
	//
	// Reference: TransformPair<T>::resonatingDispatch (TransformDSP.inl:1215-1295) feeding
	// cpl::dsp::CComplexResonator<T,2>::resonateReal, and the snapshot read by mapToLinearSpace's RSNT
	// branch (:1103-1133) through copyResonatorStateInto / getWholeWindowedState (:1152-1161).  cpl is
	// absent from the reference checkout; the filter definition is the documented decision of
	// per filter i, vector z and channel c a complex one-pole
	// s <- p s + x[n].
	//
	// Parallelisation: the recurrence is sequential in time but independent over (channel, vector,
	// filter) and over streams, so a thread owns one (c, z, i).  To fill the GPU when there are few
	// streams (BASELINE config 2 is ONE stream), the sample block is also cut into `chunks` time
	// slices: every slice runs the recurrence from a zero state in parallel (k_resonate), and
	// k_resonate_combine then chains the slices with the closed form
	//     s[n + B] = p^B s[n] + sum_j p^(B-1-j) x[n + j]
	// which is exactly what a slice computed.  Scratch: partial[stream][chunk][c][z][i].
	// ------------------------------------------------------------------------------------------
	struct ResArgs
	{
		const float* audio; uint64_t nSamples; uint64_t streamStride; int layout;
		uint64_t first;       // first sample of the block inside the buffer
		uint32_t count;       // samples in the block
		uint32_t chunks;      // time slices
		uint32_t chunkLen;    // samples per slice (last one may be shorter)
		uint32_t chunk0;      // first slice this launch computes (blockIdx.y counts from it): the tensor-core kernel takes the ones before
	};

	// channel matrix of resonatingDispatch (:1249-1294); note Mid = L + R (no 0.5) and MidSide =
	// (L - R, L + R), unlike the FFT path
	__device__ __forceinline__ float2 resonatorMatrix(int conf, float l, float r)
	{
		switch (conf)
		{
		case SGZ_CH_LEFT: return make_float2(l, 0.f);
		case SGZ_CH_RIGHT: return make_float2(r, 0.f);
		case SGZ_CH_MERGE: return make_float2(l + r, 0.f);
		case SGZ_CH_SIDE: return make_float2(l - r, 0.f);
		case SGZ_CH_MIDSIDE: return make_float2(l - r, l + r);
		default: return make_float2(l, r);
		}
	}

	constexpr int kResTile = 256;

	__global__ void __launch_bounds__(256) k_resonate(DeviceTables t, ResArgs a, float2* __restrict__ state, float2* __restrict__ partial)
	{
		__shared__ float xs[2][kResTile];
		const uint32_t P = t.P, V = t.resVectors, ch = t.resChannels;
		const uint32_t total = P * V * ch;
		const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t chunk = blockIdx.y + a.chunk0, s = blockIdx.z;
		const bool active = g < total;
		const uint32_t gi = active ? g : 0;
		const uint32_t i = gi % P, z = (gi / P) % V, c = gi / (P * V);
		const float2 p = t.resCoeff[z * P + i];
		const float* base = a.audio + uint64_t(s) * a.streamStride;

		const uint64_t begin = a.first + uint64_t(chunk) * a.chunkLen;
		const uint64_t end = min(a.first + uint64_t(a.count), begin + uint64_t(a.chunkLen));

		float2 st = make_float2(0.f, 0.f);
		if (a.chunks == 1 && active) st = state[uint64_t(s) * total + g];

		for (uint64_t n0 = begin; n0 < end; n0 += kResTile)
		{
			const uint32_t len = uint32_t(min(uint64_t(kResTile), end - n0));
			__syncthreads();
			if (threadIdx.x < len)
			{
				float l, r;
				fetchLR(base, a.layout, a.nSamples, n0 + threadIdx.x, l, r);
				const float2 x = resonatorMatrix(t.conf, l, r);
				xs[0][threadIdx.x] = x.x; xs[1][threadIdx.x] = x.y;
			}
			__syncthreads();
			const float* x = xs[c];
#pragma unroll 4
			for (uint32_t j = 0; j < len; ++j)
			{
				// s <- p s + x  (x real)
				const float re = st.x * p.x - st.y * p.y + x[j];
				const float im = st.x * p.y + st.y * p.x;
				st = make_float2(re, im);
			}
		}
		if (!active) return;
		if (a.chunks == 1) state[uint64_t(s) * total + g] = st;
		else partial[(uint64_t(s) * a.chunks + chunk) * total + g] = st;
	}

	// Two-channel form: the pole of a filter is the same for both channels, so one thread carries both states as
	// register pairs (re_1, re_2), (im_1, im_2) and every step is four packed FP32 instructions (fma.rn.f32x2 / mul)
	// for the two channels instead of eight scalar ones; the kernel is bound by instruction issue, not by memory.
	// Thread g owns (z, i) = filter g of P * V; its states sit at g (channel 1) and g + P * V (channel 2).
	__global__ void __launch_bounds__(256) k_resonate_pair(DeviceTables t, ResArgs a, float2* __restrict__ state, float2* __restrict__ partial)
	{
		__shared__ float2 xs[kResTile];
		const uint32_t P = t.P, V = t.resVectors;
		const uint32_t half = P * V, total = 2 * half;
		const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t chunk = blockIdx.y + a.chunk0, s = blockIdx.z;
		const bool active = g < half;
		const uint32_t gi = active ? g : 0;
		const float2 p = t.resCoeff[gi];   // [z][i] = gi
		const float* base = a.audio + uint64_t(s) * a.streamStride;
		const uint64_t begin = a.first + uint64_t(chunk) * a.chunkLen;
		const uint64_t end = min(a.first + uint64_t(a.count), begin + uint64_t(a.chunkLen));
		float2 re = make_float2(0.f, 0.f), im = make_float2(0.f, 0.f);
		if (a.chunks == 1 && active)
		{
			const float2 s1 = state[uint64_t(s) * total + g], s2 = state[uint64_t(s) * total + half + g];
			re = make_float2(s1.x, s2.x); im = make_float2(s1.y, s2.y);
		}
		const float2 PX = bc2(p.x), PY = bc2(p.y), NPY = bc2(-p.y);
		for (uint64_t n0 = begin; n0 < end; n0 += kResTile)
		{
			const uint32_t len = uint32_t(min(uint64_t(kResTile), end - n0));
			__syncthreads();
			if (threadIdx.x < len)
			{
				float l, r;
				fetchLR(base, a.layout, a.nSamples, n0 + threadIdx.x, l, r);
				xs[threadIdx.x] = resonatorMatrix(t.conf, l, r);
			}
			__syncthreads();
#pragma unroll 4
			for (uint32_t j = 0; j < len; ++j)
			{
				// s <- p s + x for both channels: re' = re px - im py + x, im' = re py + im px
				const float2 nre = fma2(im, NPY, fma2(re, PX, xs[j]));
				im = fma2(re, PY, mul2(im, PX));
				re = nre;
			}
		}
		if (!active) return;
		float2* dst = a.chunks == 1 ? state + uint64_t(s) * total : partial + (uint64_t(s) * a.chunks + chunk) * total;
		dst[g] = make_float2(re.x, im.x);
		dst[half + g] = make_float2(re.y, im.y);
	}

	__device__ __forceinline__ float2 cpowu(float2 p, uint32_t e)
	{
		float2 r = make_float2(1.f, 0.f);
		while (e)
		{
			if (e & 1) r = cmul(r, p);
			p = cmul(p, p);
			e >>= 1;
		}
		return r;
	}

	__global__ void __launch_bounds__(256) k_resonate_combine(DeviceTables t, ResArgs a, float2* __restrict__ state, const float2* __restrict__ partial)
	{
		const uint32_t P = t.P, V = t.resVectors, ch = t.resChannels;
		const uint32_t total = P * V * ch;
		const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t s = blockIdx.y;
		if (g >= total) return;
		const uint32_t i = g % P, z = (g / P) % V;
		const float2 p = t.resCoeff[z * P + i];
		const float2 pB = cpowu(p, a.chunkLen);
		const uint32_t lastLen = a.count - (a.chunks - 1) * a.chunkLen;
		const float2 pLast = cpowu(p, lastLen);
		float2 st = state[uint64_t(s) * total + g];
		for (uint32_t k = 0; k < a.chunks; ++k)
		{
			const float2 c = partial[(uint64_t(s) * a.chunks + k) * total + g];
			st = cadd(cmul(k + 1 == a.chunks ? pLast : pB, st), c);
		}
		state[uint64_t(s) * total + g] = st;
	}

	// getWholeWindowedState + the Phase post-combine of mapToLinearSpace's RSNT branch (:1111-1127):
	// work[x] / work[P + x] = gain_i sum_z taps_z s[c][z][i]
	__global__ void __launch_bounds__(256) k_resonator_snapshot(DeviceTables t, const float2* __restrict__ state, float2* __restrict__ workAll)
	{
		const uint32_t P = t.P, V = t.resVectors, ch = t.resChannels;
		const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
		const uint32_t s = blockIdx.y;
		if (x >= P) return;
		float2* work = workAll + uint64_t(s) * t.workStride;
		const float gain = t.resGain[x];
		float2 out[2] = { make_float2(0.f, 0.f), make_float2(0.f, 0.f) };
		for (uint32_t c = 0; c < ch; ++c)
		{
			float2 acc = make_float2(0.f, 0.f);
			for (uint32_t z = 0; z < V; ++z)
			{
				const float2 v = state[uint64_t(s) * (P * V * ch) + (c * V + z) * P + x];
				const float w = t.resTaps[z];
				acc.x += w * v.x; acc.y += w * v.y;
			}
			out[c] = make_float2(acc.x * gain, acc.y * gain);
		}
		if (t.conf == SGZ_CH_PHASE)
		{
			const float cancellation = cabs(cadd(out[0], out[1]));
			const float mid = cabs(out[0]) + cabs(out[1]);
			work[x] = make_float2(mid, 1.f - (mid > 0.f ? cancellation / mid : 0.f));
			work[P + x] = out[1];
		}
		else
		{
			work[x] = out[0];
			if (ch == 2) work[P + x] = out[1];
		}
	}
