# weak-scaling check on N GPUs of one box (N = $1), launched the way the driver launches it
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n${N}.json 2> gpurun_out/scale_n${N}.err; echo "N=$N exit $?"
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "N=1 exit $?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/scale_ref_n${N}.json 2> gpurun_out/scale_ref_n${N}.err; echo "ref N=$N exit $?"
