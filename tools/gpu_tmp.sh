for f in 0 1 2 3; do GGPM_FOLD=$f python bench.py --no-cpu-baseline --steps 5 > gpurun_out/s23_fold$f.json 2>gpurun_out/s23.err; done
