timeout 600 python bench.py --workload cfg5 > gpurun_out/s63_cfg5.json 2> gpurun_out/s63_cfg5.err; echo "cfg5 exit $?"; tail -3 gpurun_out/s63_cfg5.err
python tools/bench_summary.py gpurun_out/s63_cfg5.json
timeout 600 python bench.py --workload cfg5 --impl reference --steps 2 --warmup 1 | cut -c1-300
timeout 600 python bench.py > gpurun_out/s63_cfg2.json 2> gpurun_out/s63_cfg2.err; echo "cfg2 exit $?"; tail -3 gpurun_out/s63_cfg2.err
python tools/bench_summary.py gpurun_out/s63_cfg2.json | head -3
