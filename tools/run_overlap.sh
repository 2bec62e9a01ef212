# N=2 weak-scaling diagnosis: per-rank step times with and without the halo/compute overlap, plus the phase trace
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
python -m pytest tests/test_partition_gpu.py tests/test_apply_gpu.py tests/test_brick_gpu.py -q -m gpu -x 2>&1 | tail -3
D4B_OVERLAP_TRACE=1 $TR bench.py --gpus 2 --steps 6 --warmup 3 2>&1 >/dev/null | grep "d4b overlap" | sed -n '5,10p'
export D4B_BENCH_RANK_TIMES=1
$TR bench.py --gpus 2 --steps 50 --warmup 5 2>&1 >/dev/null | grep "\[bench\]" | sed "s/^/overlap: /"
D4B_NO_OVERLAP=1 $TR bench.py --gpus 2 --steps 50 --warmup 5 2>&1 >/dev/null | grep "\[bench\]" | sed "s/^/serial: /"
for d in 4 5; do python bench.py --deg $d --n 40 --steps 30 --no-cpu-baseline --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['config'].get('workload'), round(d['value'],2), 'GDoF/s', round(d['roofline']['frac'],3))"; done
