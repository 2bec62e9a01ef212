#!/bin/bash
# quick A/B of the hp path on the 49k-element random-hp mesh: parity of the forced two-pass path on the fixtures, then timing
cd "$(dirname "$0")/.."
python -m pytest tests/test_apply_gpu.py -q -m gpu -x -k "hp_kernel" 2>&1 | tail -2
D4B_HARNESS_NO_SOLVER=1 oracle/_ref/d4b_harness oracle/inputs/cubed_sphere_hp_l2.input random_hp 3 zero polyrand /tmp/big.fix 0 > /dev/null 2>&1
for i in 1 2; do python tools/bench_solver.py --hp /tmp/big.fix | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('two-pass', round(d['ms_per_apply'],4), 'ms', round(d['gdof_per_s'],3), 'GDoF/s rel', d['rel_l2_vs_reference_Au'])"; done
