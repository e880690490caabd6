#!/bin/bash
# Round-2 ncu evidence (one GPU, one capture per call): tools/ncu_round2.sh launches | fused | resonator
#   launches:  launch list of the bench command            -> gpurun_out/r2_launches.csv
#   fused:     full capture of the fused frame kernel      -> gpurun_out/prof_fused_r2.ncu-rep
#   resonator: full capture of the tensor-core resonator   -> gpurun_out/prof_resonator_tc_r2.ncu-rep
# each right after the same command exited 0 without ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
case "$1" in
launches)
	B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
	$B > gpurun_out/r2_plain.log 2>&1 &&
	ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/r2_ncu_l.log 2>&1
	echo "launch list rc=$?" ;;
fused)
	B2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
	$B2 > gpurun_out/r2_plain2.log 2>&1 &&
	ncu --set full --clock-control none --import-source on -k regex:k_fused8 -s 3 -c 1 -f -o gpurun_out/prof_fused_r2 $B2 > gpurun_out/r2_ncu_f.log 2>&1
	echo "fused capture rc=$?" ;;
resonator)
	R="python tools/resonator_gemm_probe.py"
	PYTHONPATH=. $R > gpurun_out/r2_plain3.log 2>&1 &&
	PYTHONPATH=. ncu --set full --clock-control none --import-source on -k regex:k_resonate_tc -s 2 -c 1 -f -o gpurun_out/prof_resonator_tc_r2 $R > gpurun_out/r2_ncu_r.log 2>&1
	echo "resonator capture rc=$?" ;;
*) echo "usage: $0 launches|fused|resonator"; exit 2 ;;
esac
