#!/bin/bash
# apply time of the hp path on small curved hp levels with and without the two-pass kernels (threshold tuning)
cd "$(dirname "$0")/.."
for f in cubed_sphere_hp1_robin cubed_sphere_hp2; do
  p=oracle/_ref/fixtures/$f.fix
  for t in 0 1000000; do
    D4B_TP_MIN=$t python tools/bench_solver.py --hp $p | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$f tp_min=$t', d['elements'], 'elements', round(d['ms_per_apply']*1000,1), 'us')"
  done
done
