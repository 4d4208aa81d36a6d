B3="python bench.py --workload cfg3s --steps 2 --warmup 3 --no-cpu-baseline --profile-step"
N="ncu --profile-from-start off --set full --clock-control none --import-source on"
$B3 > gpurun_out/s65_plain3.log 2>&1 || exit 1
python tools/bench_summary.py gpurun_out/s65_plain3.log | head -2
timeout 900 $N -k regex:"gemm_f16_kernel|lstm_sgather_kernel" -c 3 -o gpurun_out/s65_cfg3s_fwd_sgather $B3 > gpurun_out/s65_n1.log 2>&1
ls -la gpurun_out | grep s65
