T=${TAG:-cfgs}
for w in cfg1 cfg3s cfg3; do
timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; echo "$w exit $?"
python tools/bench_summary.py gpurun_out/${T}_bench_$w.json
nvidia-smi --query-gpu=memory.used --format=csv,noheader
done
