cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c5_tests.log 2>&1; tail -5 gpurun_out/r2c5_tests.log
tools/bench_variants.sh "v0 main" 24 C1,C2,C3,C4 > gpurun_out/r2c5_variants.log 2>&1; cat gpurun_out/r2c5_variants.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve -s 1 -c 1 -o gpurun_out/r2c5_solve python bench.py --nk 4 --steps 1 --warmup 1 --no-e2e --no-cpu --no-sweep > gpurun_out/r2c5_ncu.log 2>&1; tail -2 gpurun_out/r2c5_ncu.log | cut -c1-300
