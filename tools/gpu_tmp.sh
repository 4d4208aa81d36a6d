T=s46
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for r in packed make_cuda packed; do
timeout 600 python bench.py --no-cpu-baseline --resident $r > gpurun_out/${T}_b_$r.json 2> gpurun_out/${T}_b.err; echo "exit $?"; tail -2 gpurun_out/${T}_b.err
python tools/bench_summary.py gpurun_out/${T}_b_$r.json | grep mol
done
