# 8 GPUs, one box: weak scaling 1 -> 8 of cfg2, cfg5, cfg5h (quick: device-timed value only) and cfg4; one full cfg2 line at N = 8
set -x
T=${TAG:-r2n8}
run() { # name nproc workload steps extra
  if [ "$2" = "1" ]; then timeout 600 python bench.py --workload $3 --steps $4 --warmup 5 --quick $5 > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $2 --workload $3 --steps $4 --warmup 5 --quick $5 > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err; fi
  grep '^{' gpurun_out/${T}_$1.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RESULT $1', d['n_gpus'], round(d['ms_per_step'],3), round(d['value']))"; }
for W in cfg2 cfg5 cfg5h; do run ${W}_n1 1 $W 30; run ${W}_n8 8 $W 30; done
run cfg4_n1 1 cfg4 10 "--pool 4"
run cfg4_n8 8 cfg4 10 "--pool 4"
run cfg2_n8_iid 8 cfg2 30 "--no-balance"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29536 bench.py --gpus 8 --workload cfg2 --steps 20 --warmup 5 > gpurun_out/${T}_full_cfg2_n8.json 2> gpurun_out/${T}_full_cfg2_n8.err; echo "full exit $?"
python tools/bench_summary.py gpurun_out/${T}_full_cfg2_n8.json | sed -n 1,2p
