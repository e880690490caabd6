#!/bin/bash
# One GPU-box call: the Spectrum GPU tests, then bench.py; logs under gpurun_out/ (tools/gpu_check.sh [tag])
cd "$(dirname "$0")/.."
tag=${1:-check}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_spectrum_gpu.py tests/test_golden.py -x -q -m gpu > gpurun_out/${tag}_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/${tag}_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.log 2>&1
echo "bench rc=$?"; tail -1 gpurun_out/${tag}_bench.log | cut -c1-600
