# refresh the single-GPU lines of the other BASELINE configs on the round-2 build
set -x
T=${TAG:-r2g}
for W in cfg3s cfg5 cfg5h cfg3; do
timeout 1200 python bench.py --workload $W --steps 10 --warmup 3 > gpurun_out/${T}_bench_$W.json 2> gpurun_out/${T}_bench_$W.err; echo "bench $W exit $?"; tail -2 gpurun_out/${T}_bench_$W.err
python tools/bench_summary.py gpurun_out/${T}_bench_$W.json
done
