// Compiles the C++ host mirror against the C ABI and, when a GPU is present, runs one frame.
// Built and run by tests/test_cpp_mirror.py.
#include <cmath>
#include <cstdio>
#include <vector>

#include <sgz/TransformPair.hpp>

int main()
{
	sgz::TransformConstant constant;
	std::size_t transformSize = 0;
	constant.setSampleRate(48000.f);
	constant.setConfiguration(sgz::SpectrumChannels::Separate);
	constant.setBinPolation(sgz::BinInterpolation::None);
	constant.setStorage(1024, 4096, transformSize);
	constant.remapFrequencies(0.0, 1.0, sgz::ViewScaling::Logarithmic, 10.0);
	constant.generateSlopeMap(0.0, 1.0);
	constant.regenerateWindowKernel(SGZ_WIN_HANN, SGZ_WINSHAPE_PERIODIC);
	if (transformSize != 4096) { std::printf("bad transform size\n"); return 2; }
	try
	{
		sgz::TransformPairBatch pairs(constant, 2);
		const std::size_t n = 4096;
		std::vector<float> audio(2 * 2 * n); // [pair][2][n] planar
		for (std::size_t i = 0; i < n; ++i)
		{
			audio[i] = 0.5f * std::sin(6.283185307179586 * 512.0 * i / n);         // pair 0 left, bin 512
			audio[n + i] = 0.25f * std::cos(6.283185307179586 * 800.0 * i / n);    // pair 0 right
		}
		if (!pairs.prepareTransform(audio.data(), n, 2 * n)) { std::printf("unexpected skip\n"); return 3; }
		pairs.doTransform(); pairs.mapToLinearSpace(); pairs.postProcessStdTransform();
		auto res = pairs.getResults(sgz::LineMain);
		float best = -1e9f;
		for (std::size_t x = 0; x < 1024; ++x) best = std::fmax(best, res[x].leftMagnitude);
		std::printf("peak left %.5f\n", best);
		if (std::fabs(best - (1.0f - 6.0206f / 120.0f)) > 1e-4f) return 4;
		if (pairs.prepareTransform(audio.data(), n - 1, 2 * n)) return 5; // too little history -> skip
		std::printf("ok\n");
		return 0;
	}
	catch (const sgz::Error& e)
	{
		std::printf("%s\n", e.what());
		return e.code == SGZ_ECUDA ? 77 : 1; // 77: no device (expected on the CPU box)
	}
}
