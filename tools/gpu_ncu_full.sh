# ncu --set full captures of the hot kernels (one GPU call).  TAG names the outputs.
set -x
T=${TAG:-ncu}
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile-step"
N="ncu --profile-from-start off --set full --clock-control none --import-source on"
$B > gpurun_out/${T}_plain.log 2>&1 || exit 1
# cfg2, atom layer (first launches of each kernel inside the profiled step)
$N -k regex:gemm_f16_kernel -c 2 -o gpurun_out/${T}_cfg2_gemm_fwd $B > gpurun_out/${T}_n1.log 2>&1
$N -k regex:lstm_cell_fwd_kernel -s 1 -c 1 -o gpurun_out/${T}_cfg2_cell_fwd $B > gpurun_out/${T}_n2.log 2>&1
# backward runs motif -> attachment -> atom: skip the 2 x 19 small-layer rounds
$N -k regex:lstm_cell_bwd_kernel -s 40 -c 1 -o gpurun_out/${T}_cfg2_cell_bwd $B > gpurun_out/${T}_n3.log 2>&1
$N -k regex:lstm_zgather_kernel -s 40 -c 1 -o gpurun_out/${T}_cfg2_zgather $B > gpurun_out/${T}_n4.log 2>&1
$N -k regex:gemm_f16_kernel -s 137 -c 2 -o gpurun_out/${T}_cfg2_gemm_bwd $B > gpurun_out/${T}_n5.log 2>&1
# cfg3-shaped tables (beyond L2) for the HBM-side evidence
B3="python bench.py --workload cfg3s --steps 2 --warmup 3 --no-cpu-baseline --profile-step"
$B3 > gpurun_out/${T}_plain3.log 2>&1 || exit 1
timeout 900 $N -k regex:"lstm_cell_bwd_kernel|lstm_zgather_kernel" -s 80 -c 2 -o gpurun_out/${T}_cfg3s_cell_bwd_zgather $B3 > gpurun_out/${T}_n6.log 2>&1
timeout 900 $N -k regex:"lstm_cell_fwd_kernel" -s 1 -c 1 -o gpurun_out/${T}_cfg3s_cell_fwd $B3 > gpurun_out/${T}_n7.log 2>&1
timeout 900 $N -k regex:gemm_f16_kernel -s 137 -c 2 -o gpurun_out/${T}_cfg3s_gemm_bwd $B3 > gpurun_out/${T}_n8.log 2>&1
timeout 900 $N -k regex:gemm_f16_kernel -c 1 -o gpurun_out/${T}_cfg3s_gemm_fwd $B3 > gpurun_out/${T}_n9.log 2>&1
ls -la gpurun_out | grep ${T}
