# round 2, call d: whole GPU suite on the current build, cfg4 bench, A/B of the forward / state-gradient tile width, ncu of the
# backward GEMMs
set -x
T=${TAG:-r2d}
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/${T}_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench exit $?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg4.json
GGPM_F16_BN_FWD=128 timeout 900 python bench.py --no-cpu-baseline --pool 4 > gpurun_out/${T}_bench_cfg4_fwd128.json 2>/dev/null; python tools/bench_summary.py gpurun_out/${T}_bench_cfg4_fwd128.json | head -9
GGPM_F16_BN_DGRAD=128 timeout 900 python bench.py --no-cpu-baseline --pool 4 > gpurun_out/${T}_bench_cfg4_dgrad128.json 2>/dev/null; python tools/bench_summary.py gpurun_out/${T}_bench_cfg4_dgrad128.json | head -9
B4="python bench.py --workload cfg4 --steps 1 --warmup 3 --pool 2 --no-cpu-baseline --profile-step"
N="ncu --profile-from-start off --set full --clock-control none --import-source on"
timeout 500 $N -k regex:"gemm_f16_kernel" -s 200 -c 3 -o gpurun_out/${T}_cfg4_gemm_bwd $B4 > gpurun_out/${T}_n5.log 2>&1
python tools/ncu_metrics.py gpurun_out/${T}_cfg4_gemm_bwd.ncu-rep > gpurun_out/${T}_cfg4_gemm_bwd.txt 2>&1; cut -c1-160 gpurun_out/${T}_cfg4_gemm_bwd.txt | head -14
