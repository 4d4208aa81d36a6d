set -x
T=${TAG:-tmp}
timeout 600 python -m pytest tests/test_gpu_inc.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
timeout 300 python tools/inc_time.py > gpurun_out/${T}_inc_time.json 2> gpurun_out/${T}_inc_time.err; cat gpurun_out/${T}_inc_time.json
timeout 300 python tools/inc_time.py --batch 1024 --cpu-batch 0 > gpurun_out/${T}_inc_time_b1024.json 2> gpurun_out/${T}_inc_time_b1024.err; cat gpurun_out/${T}_inc_time_b1024.json
timeout 300 python tools/inc_time.py --kind motif --cpu-batch 16 > gpurun_out/${T}_inc_time_motif.json 2> gpurun_out/${T}_inc_time_motif.err; cat gpurun_out/${T}_inc_time_motif.json
