# final-build evidence (GGPM_SEG_CAP build), 1 GPU: smoke, GPU suite, reference arm, default bench (cfg4), ncu launch list of one
# cfg4 step, the other workloads' lines
set -x
T=${TAG:-r2f}
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${T}_smoke.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${T}_pytest.log
cp gpurun_out/parity_report.json gpurun_out/${T}_parity_report.json
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"; cut -c1-200 gpurun_out/${T}_bench_ref.json
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench exit $?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg4.json > gpurun_out/${T}_bench_cfg4_summary.txt; head -12 gpurun_out/${T}_bench_cfg4_summary.txt
B="python bench.py --workload cfg4 --steps 1 --warmup 3 --pool 2 --no-cpu-baseline --profile-step"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_cfg4_launches.csv $B > gpurun_out/${T}_ncu.log 2>&1; echo "ncu exit $?"
python tools/summarize_launches.py gpurun_out/${T}_cfg4_launches.csv > gpurun_out/${T}_cfg4_launches_summary.txt; head -12 gpurun_out/${T}_cfg4_launches_summary.txt
for W in cfg2 cfg1 cfg3s cfg5 cfg5h; do
timeout 600 python bench.py --workload $W --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_$W.json 2> gpurun_out/${T}_bench_$W.err; echo "$W exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_$W.json > gpurun_out/${T}_bench_${W}_summary.txt; head -1 gpurun_out/${T}_bench_${W}_summary.txt
done
