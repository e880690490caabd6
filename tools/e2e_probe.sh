#!/bin/bash
# e2e (host-buffer) throughput of bench.py for several pipeline chunk counts: tools/e2e_probe.sh 8 16 32
cd "$(dirname "$0")/.."
for c in "$@"; do
	SGZ_PIPE_CHUNKS=$c python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunks $c e2e', round(d['e2e']['value']/1e6,3), 'M frames/s', round(d['e2e']['ms_per_step'],2), 'ms')"
done
