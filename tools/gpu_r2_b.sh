# round 2, call b: GPU suite + cfg4 / cfg2 bench after the in-place backward (no dX read-modify-write)
set -x
T=${TAG:-r2b}
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/${T}_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench exit $?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg4.json
timeout 600 python bench.py --workload cfg2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; echo "cfg2 exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_cfg2.json
timeout 600 python bench.py --workload cfg1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_cfg1.json 2> gpurun_out/${T}_bench_cfg1.err; echo "cfg1 exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_cfg1.json
