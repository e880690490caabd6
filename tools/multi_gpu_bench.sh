#!/bin/bash
# bench.py on N GPUs of one box, launched like the driver launches it: tools/multi_gpu_bench.sh 2 4 8 -> gpurun_out/bench_r2_<N>gpu.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in "$@"; do
	timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 20 --warmup 3 \
		> gpurun_out/bench_r2_${n}gpu.json 2> gpurun_out/bench_r2_${n}gpu.err
	echo "N=$n rc=$?"; tail -1 gpurun_out/bench_r2_${n}gpu.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(' value', round(d['value']/1e6,2), 'M  strong_1024', round(d['strong_1024']['value']/1e6,2), 'M ratio', round(d['strong_1024']['per_gpu_rate_vs_1024_per_gpu'],3), ' gather', d['strong_1024']['gather'])
print(' e2e', round(d['e2e']['value']/1e6,3), 'M;', {k: round(v['value']/1e6,3) for k,v in d['e2e_outputs'].items()})" || tail -5 gpurun_out/bench_r2_${n}gpu.err
done
