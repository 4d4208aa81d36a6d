# ncu launch list of ONE cfg4 step (share check for the cfg4 bench line); not captured in round 1 (pod busy at round end)
T=${TAG:-cfg4l}
B="python bench.py --workload cfg4 --steps 1 --warmup 3 --no-cpu-baseline --profile-step"
timeout 200 $B > gpurun_out/${T}_plain.log 2>&1 || exit 1
timeout 200 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_launches.csv $B > gpurun_out/${T}_ncu.log 2>&1; echo "ncu exit $?"
python tools/summarize_launches.py gpurun_out/${T}_launches.csv | head -20
