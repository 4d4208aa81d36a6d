set -x
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile-step"
python -m pytest tests -m gpu -x -q > gpurun_out/s11_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s11_pytest.log
python bench.py > gpurun_out/s11_bench_cfg2.json 2> gpurun_out/s11_bench_cfg2.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s11_bench_ref.json 2>&1
python bench.py --workload cfg3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s11_bench_cfg3.json 2> gpurun_out/s11_bench_cfg3.err; echo "cfg3 exit $?"
$B > gpurun_out/s11_plain.log 2>&1 && ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s11_launches.csv $B > gpurun_out/s11_ncu1.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 40 -c 4 -o gpurun_out/s11_gemm $B > gpurun_out/s11_ncu2.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:lstm_cell -s 25 -c 2 -o gpurun_out/s11_cellfwd $B > gpurun_out/s11_ncu3.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:lstm_cell_bwd -s 5 -c 2 -o gpurun_out/s11_cellbwd $B > gpurun_out/s11_ncu4.log 2>&1
ls -la gpurun_out
