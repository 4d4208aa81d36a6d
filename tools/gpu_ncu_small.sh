B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile-step"
$B > gpurun_out/s18_plain.log 2>&1 && ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_f16_kernel -s 59 -c 4 -o gpurun_out/s18_gemm_small $B > gpurun_out/s18_ncu.log 2>&1
