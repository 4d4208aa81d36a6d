set -x
T=${TAG:-tmp}
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${T}_smoke.log
timeout 600 python -m pytest tests/test_gpu_inc.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${T}_pytest.log
timeout 300 python tools/inc_time.py --cpu-batch 0 > gpurun_out/${T}_inc_time.json 2> gpurun_out/${T}_inc_time.err; cat gpurun_out/${T}_inc_time.json
timeout 300 python tools/inc_time.py --batch 1024 --cpu-batch 0 > gpurun_out/${T}_inc_time_b1024.json 2> gpurun_out/${T}_inc_time_b1024.err; cat gpurun_out/${T}_inc_time_b1024.json
