T=s48
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
for wl in cfg2 cfg3s; do
for cfg in "0 0 0" "128 0 0" "0 128 0" "0 0 128" "128 128 128"; do
set -- $cfg
GGPM_F16_BN_FWD=$1 GGPM_F16_BN_DGRAD=$2 GGPM_F16_BN_WGRAD=$3 timeout 600 python bench.py --workload $wl --no-cpu-baseline --steps 5 > gpurun_out/${T}_b.json 2> gpurun_out/${T}_b.err
echo "$wl fwd=$1 dgrad=$2 wgrad=$3: $(python tools/bench_summary.py gpurun_out/${T}_b.json | grep 'mol\|gemm' | tr '\n' ' ')"
done
done
