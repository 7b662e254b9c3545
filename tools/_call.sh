cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q -rs ) > gpurun_out/r2af_pytest.log 2>&1; tail -4 gpurun_out/r2af_pytest.log
( time python bench.py > gpurun_out/r2af_bench_n1.json 2> gpurun_out/r2af_bench_n1.err ); tail -3 gpurun_out/r2af_bench_n1.err | cut -c1-300
B="--no-e2e --no-cpu --no-configs --no-set-c"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2af_launches.csv python bench.py --nk 16 --steps 2 --warmup 1 $B > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve -c 1 -f -o gpurun_out/r2af_solve python bench.py --nk 16 --steps 1 --warmup 1 $B --no-sweep > gpurun_out/ncu2.log 2>&1; tail -2 gpurun_out/ncu2.log
