# round 2, first GPU call: smoke, the whole -m gpu suite (no -x: measured maxima of every comparison wanted), reference
# arm, default bench (cfg4), cfg2, ncu launch list of one cfg4 step, sanitizer runs of a small step
set -x
T=${TAG:-r2a}
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${T}_smoke.log
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"; cat gpurun_out/${T}_bench_ref.json | cut -c1-600
timeout 900 python bench.py > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench exit $?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg4.json
timeout 600 python bench.py --workload cfg2 --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; echo "cfg2 exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_cfg2.json
B="python bench.py --workload cfg4 --steps 1 --warmup 3 --no-cpu-baseline --profile-step"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_cfg4_launches.csv $B > gpurun_out/${T}_ncu.log 2>&1; echo "ncu exit $?"
python tools/summarize_launches.py gpurun_out/${T}_cfg4_launches.csv > gpurun_out/${T}_cfg4_launches_summary.txt; head -25 gpurun_out/${T}_cfg4_launches_summary.txt
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_step.py > gpurun_out/${T}_sanitizer_${tool}.log 2>&1; echo "$tool exit $?"; tail -4 gpurun_out/${T}_sanitizer_${tool}.log
done
