# ncu --set full of the incremental path's kernels at 1024 molecules (largest sub-message sets), one launch each
set -x
T=${TAG:-tmp}
B="python tools/inc_time.py --batch 1024 --iters 1 --cpu-batch 0"
N="ncu --set full --clock-control none --import-source on"
$B > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err || exit 1
cat gpurun_out/${T}_plain.json
timeout 600 $N -k regex:rows_scatter_add_kernel -s 60 -c 1 -o gpurun_out/${T}_inc_scatter_add $B > gpurun_out/${T}_n1.log 2>&1
timeout 600 $N -k regex:rows_gather_kernel -s 60 -c 1 -o gpurun_out/${T}_inc_gather $B > gpurun_out/${T}_n2.log 2>&1
timeout 600 $N -k regex:sparse_lstm_cell_bwd_kernel -s 10 -c 1 -o gpurun_out/${T}_inc_cell_bwd $B > gpurun_out/${T}_n3.log 2>&1
timeout 600 $N -k regex:sparse_lstm_cell_fwd_kernel -s 10 -c 1 -o gpurun_out/${T}_inc_cell_fwd $B > gpurun_out/${T}_n4.log 2>&1
ls -la gpurun_out | grep ${T}
