# one GPU call: new-kernel tests first (bounded), then the whole GPU suite, then the benches
set -x
T=${TAG:-s12}
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "f16x3" > gpurun_out/${T}_pytest_f16.log 2>&1; echo "f16 tests exit $?" >> gpurun_out/${T}_pytest_f16.log
tail -5 gpurun_out/${T}_pytest_f16.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -5 gpurun_out/${T}_pytest.log
timeout 600 python bench.py > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; echo "bench exit $?"
timeout 600 python bench.py --engine 2 --no-cpu-baseline > gpurun_out/${T}_bench_cfg2_tf32.json 2> gpurun_out/${T}_bench_cfg2_tf32.err; echo "bench tf32 exit $?"
timeout 900 python bench.py --workload cfg3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg3.json 2> gpurun_out/${T}_bench_cfg3.err; echo "cfg3 exit $?"
