for rep in 1 2 3; do
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/tmp_b.json 2> gpurun_out/tmp_b.err; echo "exit $?"; tail -2 gpurun_out/tmp_b.err
python tools/bench_summary.py gpurun_out/tmp_b.json | grep mol
python -c "
import json; d=json.loads(open('gpurun_out/tmp_b.json').read().strip().splitlines()[-1]); print(d['clocks'])"
done
