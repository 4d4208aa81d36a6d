# what the driver runs for the scaling record at N = 8: the default bench line under torchrun, full (e2e legs, per-class pass)
set -x
T=${TAG:-r2z8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg4_n8.json 2> gpurun_out/${T}_bench_cfg4_n8.err; echo "n8 exit $?"; tail -3 gpurun_out/${T}_bench_cfg4_n8.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg4_n8.json | sed -n 1,3p
