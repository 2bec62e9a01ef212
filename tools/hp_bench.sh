#!/bin/bash
# Secondary measurement of the hp path on the GPU box: builds the 50k-element random-hp curved mesh with the unmodified
# reference + glue (no reference solve), times the apply, optionally lists the launches under ncu.
set -e
cd "$(dirname "$0")/.."
D4B_HARNESS_NO_SOLVER=1 oracle/_ref/d4b_harness oracle/inputs/cubed_sphere_hp_l2.input random_hp 3 zero polyrand /tmp/big.fix 0 > /dev/null 2>&1
python tools/bench_solver.py --hp /tmp/big.fix
if [ "$1" = "launches" ]; then
  python tools/prof_hp.py /tmp/big.fix > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:d4b_ -s 12 -c 12 --csv --log-file gpurun_out/r2_hp_launches.csv python tools/prof_hp.py /tmp/big.fix > /dev/null 2>&1
fi
