#!/bin/bash
# A/B the bench kernel over several builds made HERE (the CPU box) and shipped with the snapshot:
#   local:  tools/ab_libs.sh build t1 "-DSGZ8_STAGGER=9000" t2 "-DSGZ8_WINDOW=2" ...   (headline-only builds, ~15 s each)
#   GPU:    gpurun -- 'bash tools/ab_libs.sh run t1 t2 ...'   prints bench.py's device-resident value per build
cd "$(dirname "$0")/.."
mode=$1; shift
if [ "$mode" = build ]; then
	while [ $# -gt 1 ]; do
		make -C signalizer_b200/csrc -s -j4 VARIANT=$1 DEFS="-DSGZ_HEADLINE_ONLY $2" || exit 1
		echo -n "$1 [$2]: "; grep -A3 "k_fused8ILi8ELb1ELi2ELi0" signalizer_b200/csrc/build_$1/sgz_spectrum.log | grep -E "spill|Used" | tr '\n' ' '; echo
		shift 2
	done
else
	for v in "$@"; do
		for i in 1 2; do
			SGZ_CUDA_LIB=$PWD/signalizer_b200/libsgz_cuda_$v.so python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e 2>gpurun_out/ab_$v.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', 'value', round(d['value']/1e6,3), 'M frames/s  frac', round(d['roofline']['frac'],4), 'clk', d['clocks']['sm_mhz'])" || tail -3 gpurun_out/ab_$v.err
		done
	done
fi
