T=${TAG:-tmp}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
for W in cfg2 cfg1 cfg3s; do
  timeout 600 python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_${W}.json 2> gpurun_out/${T}_${W}.err
  python tools/bench_summary.py gpurun_out/${T}_${W}.json | head -2
done
