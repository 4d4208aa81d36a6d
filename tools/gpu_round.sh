# what the driver runs at round end, in one call: build check, smoke, GPU tests, reference arm, our arm; then cfg3
set -x
T=${TAG:-round}
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/${T}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"
timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"; tail -2 gpurun_out/${T}_bench.err
python tools/bench_summary.py gpurun_out/${T}_bench.json
timeout 1200 python bench.py --workload cfg3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg3.json 2> gpurun_out/${T}_bench_cfg3.err; echo "cfg3 exit $?"; tail -2 gpurun_out/${T}_bench_cfg3.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg3.json
