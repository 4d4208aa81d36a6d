set -x
T=${TAG:-r2f}
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/${T}_pytest.log
for W in cfg4 cfg2 cfg1; do
timeout 900 python bench.py --workload $W --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_$W.json 2> gpurun_out/${T}_bench_$W.err; echo "bench $W exit $?"; tail -2 gpurun_out/${T}_bench_$W.err
python tools/bench_summary.py gpurun_out/${T}_bench_$W.json
done
