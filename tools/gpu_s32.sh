set -x
T=s33
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; echo "bench exit $?"
GGPM_PDL=0 timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench_cfg2_nopdl.json 2> gpurun_out/${T}_bench_cfg2_nopdl.err; echo "bench exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_cfg2.json gpurun_out/${T}_bench_cfg2_nopdl.json
