#!/bin/bash
# A/B compile-time switches of the CUDA library against a probe script: tools/ab_probe.sh tools/rs_probe.py "-DX=1" "-DX=2"
cd "$(dirname "$0")/.."
probe=$1; shift
for flags in "$@"; do
	touch signalizer_b200/csrc/sgz_scope.cu signalizer_b200/csrc/sgz_spectrum.cu
	make -C signalizer_b200/csrc -s NVFLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v --resource-usage $flags" > /dev/null 2>&1
	echo "== $flags"
	PYTHONPATH=. python $probe 2>&1 | tail -4
done
