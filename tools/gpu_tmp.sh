T=${TAG:-tmp}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
for OV in 0 1; do
for W in cfg2 cfg3s; do
  GGPM_WGRAD_OVERLAP=$OV timeout 600 python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_${W}_ov${OV}.json 2> gpurun_out/${T}_${W}_ov${OV}.err
  echo "overlap=$OV"; python tools/bench_summary.py gpurun_out/${T}_${W}_ov${OV}.json | head -2
done
done
timeout 900 python bench.py --workload cfg4 --steps 32 --warmup 4 --no-cpu-baseline > gpurun_out/${T}_cfg4.json 2> gpurun_out/${T}_cfg4.err; echo "cfg4 exit $?"; tail -2 gpurun_out/${T}_cfg4.err
python tools/bench_summary.py gpurun_out/${T}_cfg4.json | head -2
