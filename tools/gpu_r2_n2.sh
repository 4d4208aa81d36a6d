# 2 GPUs: NCCL parity test of the grouped / overlapped gradient reduction, then cfg2 and cfg5 bench lines at N = 2
set -x
T=${TAG:-r2n2}
timeout 900 python -m pytest tests/test_gpu_nccl.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${T}_pytest.log
cat gpurun_out/nccl_parity_*.json
for W in cfg2 cfg5; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --workload $W --steps 20 --warmup 5 > gpurun_out/${T}_bench_${W}.json 2> gpurun_out/${T}_bench_${W}.err; echo "$W exit $?"; tail -2 gpurun_out/${T}_bench_${W}.err
  python tools/bench_summary.py gpurun_out/${T}_bench_${W}.json | head -2
done
