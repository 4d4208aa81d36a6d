# 2 GPUs, final build: NCCL parity test of the gradient reduction, the default bench line and the reference arm at N = 2
set -x
T=${TAG:-r2n2d}
timeout 600 python -m pytest tests/test_gpu_nccl.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${T}_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 4 --no-cpu-baseline > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "cfg4 exit $?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg4.json | head -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"; cut -c1-160 gpurun_out/${T}_bench_ref.json
