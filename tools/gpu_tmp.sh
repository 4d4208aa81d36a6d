set -x
T=${TAG:-tmp}
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/${T}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
timeout 300 python tools/inc_time.py > gpurun_out/${T}_inc_time.json 2> gpurun_out/${T}_inc_time.err; cat gpurun_out/${T}_inc_time.json
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"; tail -2 gpurun_out/${T}_bench.err
python tools/bench_summary.py gpurun_out/${T}_bench.json
