cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q -rs ) > gpurun_out/r2h_pytest.log 2>&1; grep -E "passed|failed|FAILED" gpurun_out/r2h_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-250
( time python bench.py > gpurun_out/r2h_bench_n1.json 2> gpurun_out/r2h_bench_n1.err ); tail -2 gpurun_out/r2h_bench_n1.err | cut -c1-300
