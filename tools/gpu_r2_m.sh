set -x
T=${TAG:-r2m}
timeout 1200 python -m pytest tests/test_gpu_encoder.py -m gpu -q -x -k "epilogue or cfg1_matches or plane_fed" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/${T}_pytest.log | cut -c1-200
GGPM_FUSED_GATES=1 timeout 900 python -m pytest tests/test_gpu_parity_full.py -m gpu -q -x -k "plane_fed or hidden_500_depth" > gpurun_out/${T}_pytest2.log 2>&1; echo "pytest2 exit $?"; tail -5 gpurun_out/${T}_pytest2.log | cut -c1-200
for F in 0 1 0 1; do
GGPM_FUSED_GATES=$F timeout 600 python bench.py --workload cfg4 --steps 20 --warmup 5 --pool 4 --no-cpu-baseline > gpurun_out/${T}_fused$F.json 2>/dev/null
echo "FUSED $F"; python tools/bench_summary.py gpurun_out/${T}_fused$F.json | grep -v roofline
done
