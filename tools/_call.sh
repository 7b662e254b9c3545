cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -5
python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_parity_large.py 2>&1 | tail -5
