# ncu --set full of the row kernels at cfg4 (batch 4096, hidden 250: atom layer E+1 ~ 306 k rows x Hp = 256, tables beyond the L2)
set -x
T=${TAG:-ncu4}
B4="python bench.py --workload cfg4 --steps 1 --warmup 3 --no-cpu-baseline --profile-step"
N="ncu --profile-from-start off --set full --clock-control none --import-source on"
timeout 300 $B4 > gpurun_out/${T}_plain.log 2>&1 || exit 1
# backward runs motif -> attachment -> atom: skip the 2 x 19 small-layer rounds of each kernel
timeout 400 $N -k regex:"lstm_cell_bwd_kernel|lstm_zgather_kernel" -s 80 -c 2 -o gpurun_out/${T}_cfg4_cell_bwd_zgather $B4 > gpurun_out/${T}_n1.log 2>&1
timeout 400 $N -k regex:"lstm_cell_fwd_kernel" -s 1 -c 1 -o gpurun_out/${T}_cfg4_cell_fwd $B4 > gpurun_out/${T}_n2.log 2>&1
ls -la gpurun_out | grep ${T}
