# ncu --set full of the top kernel (backward cell kernel, atom layer) and of the two backward GEMMs of one cfg4 step, final build
set -x
T=${TAG:-r2ncuf}
B4="python bench.py --workload cfg4 --steps 1 --warmup 3 --pool 2 --no-cpu-baseline --profile-step"
N="ncu --profile-from-start off --set full --clock-control none --import-source on"
timeout 300 $B4 > gpurun_out/${T}_plain.log 2>&1 || exit 1
timeout 400 $N -k regex:"lstm_cell_bwd_kernel" -s 45 -c 1 -o gpurun_out/${T}_cfg4_cell_bwd $B4 > gpurun_out/${T}_n1.log 2>&1
timeout 400 $N -k regex:"gemm_f16_kernel" -s 250 -c 2 -o gpurun_out/${T}_cfg4_gemm_bwd $B4 > gpurun_out/${T}_n5.log 2>&1
for f in gpurun_out/${T}_cfg4_*.ncu-rep; do python tools/ncu_metrics.py $f > ${f%.ncu-rep}.txt 2>&1; done
head -40 gpurun_out/${T}_cfg4_cell_bwd.txt
