set -x
T=s30
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 600 python bench.py > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"
timeout 300 python tools/f16_gemm_time.py > gpurun_out/${T}_f16_gemm_time.txt 2>&1
timeout 300 python tools/f16_trace.py > gpurun_out/${T}_f16_trace.txt 2>&1
timeout 300 python tools/host_time.py > gpurun_out/${T}_host_time.txt 2>&1
cat gpurun_out/${T}_f16_gemm_time.txt gpurun_out/${T}_f16_trace.txt
tail -3 gpurun_out/${T}_host_time.txt
