set -x
T=${TAG:-r2l}
timeout 600 python bench.py --workload cfg2g --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg2g.json 2> gpurun_out/${T}_bench_cfg2g.err; echo "exit $?"; tail -3 gpurun_out/${T}_bench_cfg2g.err
python tools/bench_summary.py gpurun_out/${T}_bench_cfg2g.json
timeout 300 python tools/step_timeline.py --workload cfg2g > gpurun_out/${T}_timeline_cfg2g.txt 2>/dev/null; head -24 gpurun_out/${T}_timeline_cfg2g.txt
