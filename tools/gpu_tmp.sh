timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/s51_n2.json 2> gpurun_out/s51_n2.err; echo "N=2 exit $?"; tail -2 gpurun_out/s51_n2.err
python tools/bench_summary.py gpurun_out/s51_n2.json | grep mol
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/s51_ref_n2.json 2> gpurun_out/s51_ref_n2.err; echo "ref N=2 exit $?"
tail -1 gpurun_out/s51_ref_n2.json | cut -c1-200
timeout 600 python bench.py > gpurun_out/s51_n1.json 2> gpurun_out/s51_n1.err; echo "N=1 exit $?"
python tools/bench_summary.py gpurun_out/s51_n1.json | grep mol
