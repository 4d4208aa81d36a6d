T=${TAG:-cfgs}
for w in ${WL:-cfg3 cfg2}; do
timeout 1200 python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; echo "$w exit $?"
tail -2 gpurun_out/${T}_bench_$w.err
python tools/bench_summary.py gpurun_out/${T}_bench_$w.json
done
