# one GPU call: the GPU test suite, then the default bench; TAG names the outputs
T=${TAG:-run}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
timeout 600 python bench.py ${BENCH_ARGS:---no-cpu-baseline} > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"
tail -3 gpurun_out/${T}_bench.err
python tools/bench_summary.py gpurun_out/${T}_bench.json
python -c "
import json; d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1]); print({k:(round(v['value']),round(v['ms_per_step'],2),v['h2d_bytes_per_step']) for k,v in d.items() if k.startswith('e2e')})"
