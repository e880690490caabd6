#!/bin/bash
# A/B a compile-time switch of the CUDA library on the GPU box: tools/ab_bench.sh "-DSGZ8_TW_POWERS=0" "-DSGZ8_TW_POWERS=1"
# rebuilds libsgz_cuda.so with each flag set in turn and prints bench.py's value for it (twice each).
cd "$(dirname "$0")/.."
for flags in "$@"; do
	touch signalizer_b200/csrc/sgz_spectrum.cu
	make -C signalizer_b200/csrc -s NVFLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v --resource-usage $flags" > /dev/null 2>&1
	grep -A3 "k_fused8ILi8ELb1ELi2ELi0" signalizer_b200/csrc/build.log | grep -E "spill|Used" | tr '\n' ' '
	echo
	for i in 1 2; do
		python bench.py --steps 30 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$flags', 'value', round(d['value']/1e6,3), 'M frames/s  frac', round(d['roofline']['frac'],4), 'clk', d['clocks']['sm_mhz'])"
	done
done
