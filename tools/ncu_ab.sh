#!/bin/bash
# ncu --set full capture of the fused kernel for A/B builds: tools/ncu_ab.sh old w3 ...  -> gpurun_out/prof_<v>.ncu-rep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in "$@"; do
	export SGZ_CUDA_LIB=$PWD/signalizer_b200/libsgz_cuda_$v.so
	python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain_$v.log 2>&1 &&
	ncu --set full --clock-control none --import-source on -k regex:k_fused8 -s 3 -c 1 -f -o gpurun_out/prof_$v python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_$v.log 2>&1
	echo "$v rc=$?"
done
