// sgz_harness.cpp — the batched offline harness of the Spectrum hot path in C++ (north_star: "Host code is C++ calling a
// thin C-ABI layer"; SURVEY.md §7 step 5: one host thread + one handle + one stream per device).
//
// What the reference does per audio callback for at most eight stereo pairs - AudioDispatcher::dispatch looping
// parallel_for over access->pairs, each pair running its frames (Source/Spectrum/SpectrumDSP.cpp:61-108) - this program
// does for a wall of independent streams over every GPU of the box: streams are sharded contiguously over the devices
// (the same rule as signalizer_b200/sharding.py), every device gets its own std::thread, sgz_spectrum handle and
// device-resident buffers, nothing is exchanged between devices.  It prints one JSON line with the aggregate frames/s
// (device time: cudaEvents on each handle's stream, the slowest device decides) and a checksum of the last frame.
//
//   g++ -std=c++17 -O2 -Iinclude -I/usr/local/cuda/include tools/sgz_harness.cpp -Lsignalizer_b200 -lsgz_cuda \
//       -L/usr/local/cuda/lib64 -lcudart -lpthread -Wl,-rpath,$PWD/signalizer_b200 -o sgz_harness
//   ./sgz_harness [streams=1024] [frames=64] [steps=20] [devices=all]
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include <sgz.h>

namespace
{
	// the synthetic signals of SURVEY.md §8(d), sine / chirp / noise by stream index (signalizer_b200/signals.py)
	uint64_t splitmix64(uint64_t x)
	{
		x += 0x9E3779B97F4A7C15ull;
		x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
		x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
		return x ^ (x >> 31);
	}

	void synth(float* dst, uint32_t stream, uint64_t n, double fs)
	{
		const double pi = 3.14159265358979323846;
		for (uint64_t i = 0; i < n; ++i)
		{
			float l, r;
			switch (stream % 3)
			{
			case 0: { const double f = 55.0 * std::pow(2.0, double(stream % 96) / 12.0); l = float(0.5 * std::sin(2 * pi * f * double(i) / fs)); r = float(0.25 * std::sin(2 * pi * f * double(i) / fs + pi / 3)); break; }
			case 1: { const double T = double(n) / fs, k = std::log(20000.0 / 20.0) / T, t = double(i) / fs; const double ph = 2 * pi * 20.0 * (std::exp(k * t) - 1.0) / k; l = float(0.5 * std::sin(ph)); r = float(0.5 * std::cos(ph)); break; }
			default: {
				const uint64_t a = splitmix64(0x5167A11Eull ^ (uint64_t(stream) << 33) ^ (0ull << 32) ^ i) >> 40, b = splitmix64(0x5167A11Eull ^ (uint64_t(stream) << 33) ^ (1ull << 32) ^ i) >> 40;
				l = float(double(a) * (1.0 / 8388608.0) - 1.0); r = float(double(b) * (1.0 / 8388608.0) - 1.0); break; }
			}
			dst[2 * i] = l; dst[2 * i + 1] = r;
		}
	}

	struct DeviceJob
	{
		int device; uint32_t first, count; // streams [first, first + count)
		double ms{ 0 }; double checksum{ 0 }; std::string error;
	};

	void run(DeviceJob& job, const sgz_spectrum_config& cfg, uint32_t frames, uint32_t hop, uint32_t steps)
	{
		auto failed = [&](const char* what) { job.error = std::string(what) + ": " + sgz_last_error(); };
		if (cudaSetDevice(job.device) != cudaSuccess) { job.error = "cudaSetDevice failed"; return; }
		sgz_spectrum* h = nullptr;
		if (sgz_spectrum_create(&cfg, job.count, job.device, &h) != SGZ_OK) return failed("sgz_spectrum_create");
		const uint64_t n = uint64_t(cfg.window_size) + uint64_t(frames - 1) * hop, stride = 2 * n;
		const uint32_t P = cfg.axis_points;
		std::vector<float> host(size_t(job.count) * stride);
		for (uint32_t s = 0; s < job.count; ++s) synth(host.data() + size_t(s) * stride, (job.first + s) % 96, n, cfg.sample_rate);
		float *dAudio = nullptr, *dOut = nullptr;
		const size_t outFloats = size_t(job.count) * frames * 2 * P * 2;
		cudaStream_t stream = nullptr;
		cudaEvent_t e0 = nullptr, e1 = nullptr;
		if (cudaMalloc(&dAudio, host.size() * sizeof(float)) != cudaSuccess || cudaMalloc(&dOut, outFloats * sizeof(float)) != cudaSuccess ||
			cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess)
		{ job.error = "device allocation failed"; sgz_spectrum_destroy(h); return; }
		cudaMemcpy(dAudio, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice);
		sgz_spectrum_set_cuda_stream(h, stream);
		auto step = [&]() { return sgz_spectrum_process_frames(h, dAudio, SGZ_MEM_DEVICE, SGZ_LAYOUT_INTERLEAVED, n, stride, hop, frames, dOut, SGZ_MEM_DEVICE); };
		bool ok = true;
		for (int w = 0; w < 3 && ok; ++w) ok = step() == SGZ_OK;
		cudaStreamSynchronize(stream);
		cudaEventRecord(e0, stream);
		for (uint32_t i = 0; i < steps && ok; ++i) ok = step() == SGZ_OK;
		cudaEventRecord(e1, stream);
		cudaStreamSynchronize(stream);
		if (!ok) failed("sgz_spectrum_process_frames");
		else
		{
			float ms = 0; cudaEventElapsedTime(&ms, e0, e1); job.ms = ms / steps;
			std::vector<float> last(size_t(2) * P * 2);
			double sum = 0;
			for (uint32_t s = 0; s < job.count; s += std::max<uint32_t>(1, job.count / 16))
			{
				cudaMemcpy(last.data(), dOut + (size_t(s) * frames + (frames - 1)) * 2 * P * 2, last.size() * sizeof(float), cudaMemcpyDeviceToHost);
				for (uint32_t x = 0; x < P; ++x) sum += last[2 * x];
			}
			job.checksum = sum;
		}
		cudaFree(dAudio); cudaFree(dOut); cudaEventDestroy(e0); cudaEventDestroy(e1);
		sgz_spectrum_destroy(h);
		cudaStreamDestroy(stream);
	}
}

int main(int argc, char** argv)
{
	const uint32_t streams = argc > 1 ? uint32_t(std::atoi(argv[1])) : 1024, frames = argc > 2 ? uint32_t(std::atoi(argv[2])) : 64;
	const uint32_t steps = argc > 3 ? uint32_t(std::atoi(argv[3])) : 20;
	int devices = 0;
	if (cudaGetDeviceCount(&devices) != cudaSuccess || devices == 0) { std::printf("{\"error\": \"no CUDA device (there is no CPU fallback)\"}\n"); return 77; }
	if (argc > 4) devices = std::min(devices, std::atoi(argv[4]));
	devices = int(std::min<uint32_t>(uint32_t(devices), streams));
	sgz_spectrum_config cfg;
	sgz_spectrum_default_config(&cfg); // BASELINE config 3: 8192-pt Blackman-Harris, Separate, Lanczos, 1024 log bins
	const uint32_t hop = cfg.window_size / 4;
	std::vector<DeviceJob> jobs(devices);
	const uint32_t base = streams / devices, extra = streams % devices;
	for (int d = 0; d < devices; ++d)
	{
		jobs[d].device = d;
		jobs[d].first = uint32_t(d) * base + std::min<uint32_t>(uint32_t(d), extra);
		jobs[d].count = base + (uint32_t(d) < extra ? 1u : 0u);
	}
	std::vector<std::thread> pool;
	for (auto& j : jobs) pool.emplace_back(run, std::ref(j), std::cref(cfg), frames, hop, steps);
	for (auto& t : pool) t.join();
	double slowest = 0, checksum = 0;
	for (const auto& j : jobs)
	{
		if (!j.error.empty()) { std::printf("{\"error\": \"device %d: %s\"}\n", j.device, j.error.c_str()); return 1; }
		slowest = std::max(slowest, j.ms); checksum += j.checksum;
	}
	std::printf("{\"harness\": \"sgz_harness (C++)\", \"devices\": %d, \"streams\": %u, \"frames_per_stream\": %u, \"steps\": %u, \"ms_per_step\": %.4f, "
		"\"frames_per_s\": %.1f, \"checksum\": %.6f}\n", devices, streams, frames, steps, slowest, double(streams) * frames / (slowest * 1e-3), checksum);
	return 0;
}
