# ncu --set full of the dominant kernels of one cfg4 step (batch 4096, hidden 250: atom layer 306 k rows x Hp 256, tables
# beyond the L2), after the program exits 0 without ncu.  One step runs forward atom -> attachment -> motif, backward in
# reverse; -s skips the launches of the two small layers so that each capture is an ATOM-layer launch.
set -x
T=${TAG:-r2ncu}
B4="python bench.py --workload cfg4 --steps 1 --warmup 3 --pool 2 --no-cpu-baseline --profile-step"
N="ncu --profile-from-start off --set full --clock-control none --import-source on"
timeout 300 $B4 > gpurun_out/${T}_plain.log 2>&1 || exit 1
timeout 500 $N -k regex:"lstm_cell_bwd_kernel" -s 45 -c 1 -o gpurun_out/${T}_cfg4_cell_bwd $B4 > gpurun_out/${T}_n1.log 2>&1
timeout 500 $N -k regex:"lstm_zgather_kernel" -s 42 -c 1 -o gpurun_out/${T}_cfg4_zgather $B4 > gpurun_out/${T}_n2.log 2>&1
timeout 500 $N -k regex:"lstm_cell_fwd_kernel" -s 3 -c 1 -o gpurun_out/${T}_cfg4_cell_fwd $B4 > gpurun_out/${T}_n3.log 2>&1
timeout 500 $N -k regex:"gemm_f16_kernel" -s 4 -c 2 -o gpurun_out/${T}_cfg4_gemm_fwd $B4 > gpurun_out/${T}_n4.log 2>&1
timeout 500 $N -k regex:"gemm_f16_kernel" -s 250 -c 2 -o gpurun_out/${T}_cfg4_gemm_bwd $B4 > gpurun_out/${T}_n5.log 2>&1
timeout 500 $N -k regex:"lstm_sgather_kernel|lstm_dx_reduce_kernel" -s 3 -c 1 -o gpurun_out/${T}_cfg4_sgather $B4 > gpurun_out/${T}_n6.log 2>&1
ls -la gpurun_out | grep ${T}
for f in gpurun_out/${T}_cfg4_*.ncu-rep; do python tools/ncu_metrics.py $f > ${f%.ncu-rep}.txt 2>&1; done
