set -x
T=${TAG:-r2e}
timeout 600 python tools/step_timeline.py --workload cfg4 > gpurun_out/${T}_timeline_cfg4.txt 2> gpurun_out/${T}_timeline_cfg4.err; tail -3 gpurun_out/${T}_timeline_cfg4.err; cat gpurun_out/${T}_timeline_cfg4.txt
timeout 600 python tools/step_timeline.py --workload cfg2 > gpurun_out/${T}_timeline_cfg2.txt 2> gpurun_out/${T}_timeline_cfg2.err; cat gpurun_out/${T}_timeline_cfg2.txt
