for wl in cfg3s; do
for bn in 0 128 0 128; do
GGPM_F16_BN_FWDP=$bn timeout 600 python bench.py --workload $wl --no-cpu-baseline --steps 10 > gpurun_out/s70_b.json 2> gpurun_out/s70_b.err; tail -1 gpurun_out/s70_b.err
echo "$wl fwdp=$bn: $(python tools/bench_summary.py gpurun_out/s70_b.json | grep 'mol\|gemm_tcgen05 ' | tr '\n' ' ' | cut -c1-300)"
done
done
