cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python tools/calc_probe.py 128 > gpurun_out/r2ag_calc.txt 2>&1; cat gpurun_out/r2ag_calc.txt
python -m pytest tests/test_gpu_parity.py tests/test_gpu_reference_suite.py -m gpu -x -q 2>&1 | tail -3
