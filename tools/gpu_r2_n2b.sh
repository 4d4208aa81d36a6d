# same box: cfg2 at N = 1, N = 2 with the overlapped grouped reduction, N = 2 with one group after the backward pass
set -x
T=${TAG:-r2n2b}
run() { # name, nproc, extra env
  if [ "$2" = "1" ]; then env $3 timeout 300 python bench.py --workload $4 --steps 30 --warmup 5 --quick > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err
  else env $3 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $2 --workload $4 --steps 30 --warmup 5 --quick > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err; fi
  echo "$1: $(cut -c1-230 gpurun_out/${T}_$1.json)"; }
run cfg2_n1 1 X=1 cfg2
run cfg2_n2_overlap 2 GGPM_BUCKET_OVERLAP=1 cfg2
run cfg2_n2_single 2 GGPM_BUCKET_OVERLAP=0 cfg2
run cfg2_n1_again 1 X=1 cfg2
run cfg5h_n1 1 X=1 cfg5h
run cfg5h_n2 2 X=1 cfg5h
