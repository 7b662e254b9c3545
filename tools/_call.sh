cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CONFIGS=C3 python tools/bench_configs.py 23 B 2>&1 | grep -E "Mcell|rror"
CONFIGS=HARNESS python tools/bench_configs.py 21 H 2>&1 | grep -E "Mcell|rror"
python -m pytest tests/test_gpu_parity.py tests/test_gpu_variants.py tests/test_gpu_parity_large.py -m gpu -x -q 2>&1 | tail -3
