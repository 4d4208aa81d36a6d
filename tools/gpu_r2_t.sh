# accumulation-chain cap (GGPM_SEG_CAP) validation, 1 GPU: the allocation-free entry point alone in a fresh process, the two
# sweep-size parity tests at the default cap, the 4096-molecule one again with the cap off (expected to FAIL: that is the
# A/B), cfg4 / cfg2 step time at both settings, then the whole GPU suite
set -x
T=${TAG:-r2t}
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${T}_noalloc_alone.log 2>&1; echo "noalloc-alone exit $?"; tail -3 gpurun_out/${T}_noalloc_alone.log
timeout 900 python -m pytest tests/test_gpu_parity_full.py -m gpu -q -k evaluated_on_the_gpu > gpurun_out/${T}_full_cap16.log 2>&1; echo "cap16 exit $?"; tail -4 gpurun_out/${T}_full_cap16.log
cp gpurun_out/parity_report.json gpurun_out/${T}_parity_cap16.json
GGPM_SEG_CAP=0 timeout 600 python -m pytest tests/test_gpu_parity_full.py -m gpu -q -k sweep_batch_of_4096 > gpurun_out/${T}_full_cap0.log 2>&1; echo "cap0 exit $? (failure expected)"; tail -4 gpurun_out/${T}_full_cap0.log
cp gpurun_out/parity_report.json gpurun_out/${T}_parity_cap0.json
for CAP in 0 16; do
for W in cfg4 cfg2; do
GGPM_SEG_CAP=$CAP timeout 600 python bench.py --workload $W --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_${W}_cap$CAP.json 2> gpurun_out/${T}_bench_${W}_cap$CAP.err; echo "$W cap $CAP exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_${W}_cap$CAP.json | head -2
done
done
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/${T}_pytest.log
cp gpurun_out/parity_report.json gpurun_out/${T}_parity_report.json
