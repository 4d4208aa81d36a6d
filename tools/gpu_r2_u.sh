# hidden 500 x 1024 molecules: where the gradient error comes from (per engine, on the product's ReLU pattern), then the two
# sweep-size parity tests with the fp32-on-the-same-branch tolerance
set -x
T=${TAG:-r2u}
timeout 600 python tools/diag_precision.py --hidden 500 --batch 1024 --seed 31024 > gpurun_out/${T}_diag_h500.txt 2>&1; echo "diag exit $?"; tail -45 gpurun_out/${T}_diag_h500.txt
timeout 900 python -m pytest tests/test_gpu_parity_full.py -m gpu -q -k evaluated_on_the_gpu > gpurun_out/${T}_full.log 2>&1; echo "full exit $?"; tail -6 gpurun_out/${T}_full.log
cp gpurun_out/parity_report.json gpurun_out/${T}_parity_report.json
