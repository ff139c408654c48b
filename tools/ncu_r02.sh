#!/bin/bash
# Round-2 ncu captures (run on the GPU box through gpurun; every bench command below has exited 0 without ncu before):
#   launch list of the default bench command, then `--set full` captures of the dominant kernel(s) of every configuration.
set -x
O=gpurun_out
NCU="ncu --clock-control none"
$NCU --metrics gpu__time_duration.sum -c 900 --csv --log-file $O/r02_launches_default.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --fb-utts 100000 > $O/ncu_launches.log 2>&1
full() { name=$1; shift; $NCU --set full --import-source on "$@" > $O/ncu_$name.log 2>&1; ncu -i $O/r02_$name.ncu-rep --page raw --csv > $O/r02_ncu_${name}_raw.csv 2>/dev/null; }
full em      -k regex:"em_isolated|em_push|em_update" -c 3 -o $O/r02_em python bench.py --workload em --no-sub --no-cpu --no-e2e --steps 1 --warmup 0
full decode  -k regex:classify_tc -c 1 -o $O/r02_decode python bench.py --workload decode --no-cpu --no-e2e --steps 1 --warmup 0
full stress  -k regex:gmm_best_gemm -c 1 -o $O/r02_stress python bench.py --workload stress --no-cpu --no-e2e --steps 1 --warmup 0
full connected -k regex:"viterbi_banded1|classify_tc" -c 2 -o $O/r02_connected python bench.py --workload connected --no-cpu --no-e2e --steps 1 --warmup 0
full fb      -k regex:"gmm_stats_soft_reg|forward_backward_banded|gmm_best_gemm" -c 3 -o $O/r02_fb python bench.py --workload fb --fb-utts 200000 --no-cpu --no-e2e --steps 1 --warmup 0
ls -la $O/*.ncu-rep $O/r02_ncu_*_raw.csv
