# weak-scaling check on the GPUs of one box (N = $1), launched the way the driver launches it: N = 1, 2, 4 .. $1
NMAX=${1:-2}
for N in 1 2 4 8; do
  [ $N -gt $NMAX ] && break
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/scale_n${N}.json 2> gpurun_out/scale_n${N}.err
  fi
  echo "N=$N exit $?: $(python tools/bench_summary.py gpurun_out/scale_n${N}.json | grep mol)"
done
