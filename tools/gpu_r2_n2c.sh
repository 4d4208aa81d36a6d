set -x
T=${TAG:-r2n2c}
run() { env $3 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $2 --workload $4 --steps 30 --warmup 5 --quick > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err; grep '^{' gpurun_out/${T}_$1.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['n_gpus'], round(d['ms_per_step'],3))"; }
timeout 300 python bench.py --workload cfg2 --steps 30 --warmup 5 --quick | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('n1', round(d['ms_per_step'],3))"
run n2_default 2 X=1 cfg2
run n2_same_seed 2 GGPM_BENCH_SAME_SEED=1 cfg2
run n2_no_reduce 2 GGPM_BENCH_NO_REDUCE=1 cfg2
run n2_same_seed_no_reduce 2 "GGPM_BENCH_SAME_SEED=1 GGPM_BENCH_NO_REDUCE=1" cfg2
