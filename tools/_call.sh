cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q -rs ) > gpurun_out/r2f_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/r2f_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-250
( time python bench.py > gpurun_out/r2f_bench_n1.json 2> gpurun_out/r2f_bench_n1.err ); tail -2 gpurun_out/r2f_bench_n1.err | cut -c1-300
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2f_bench_ref.json 2> gpurun_out/r2f_bench_ref.err )
B="--no-e2e --no-cpu --no-configs --no-set-c"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2f_launches.csv python bench.py --nk 16 --steps 2 --warmup 1 $B > gpurun_out/ncu1.log 2>&1
