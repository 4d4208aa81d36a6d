# BASELINE config #4: 1M synthetic molecular graphs in batches of 4096 per GPU (pool of 16 distinct batches rotating)
# N=1: 245 steps; under torchrun with N ranks: ceil(245 / N) steps per rank
T=${TAG:-cfg4}
N=${NGPU:-1}
STEPS=$(( (245 + N - 1) / N ))
if [ "$N" = "1" ]; then
  timeout 1200 python bench.py --workload cfg4 --steps $STEPS --warmup 4 --no-cpu-baseline > gpurun_out/${T}_n1.json 2> gpurun_out/${T}_n1.err
else
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --workload cfg4 --steps $STEPS --warmup 4 --no-cpu-baseline > gpurun_out/${T}_n${N}.json 2> gpurun_out/${T}_n${N}.err
fi
echo "cfg4 N=$N exit $?"; tail -2 gpurun_out/${T}_n${N}.err
python tools/bench_summary.py gpurun_out/${T}_n${N}.json | head -2
