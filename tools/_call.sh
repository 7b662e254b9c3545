cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out; rm -f gpurun_out/parity_large.txt
( time python -m pytest tests -m gpu -q --durations=6 ) > gpurun_out/r2d1_tests.log 2>&1; tail -45 gpurun_out/r2d1_tests.log | cut -c1-700
grep -v calculate_ gpurun_out/parity_large.txt
