cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_multi_device.py tests/test_zz_no_alloc.py tests/test_gpu_flat_api.py -m gpu -x -q 2>&1 | tail -4
python tools/concurrent_callers_probe.py 8 20 > gpurun_out/r2ac_concurrent.txt 2>&1
python tools/concurrent_callers_probe.py 16 18 >> gpurun_out/r2ac_concurrent.txt 2>&1
cat gpurun_out/r2ac_concurrent.txt | cut -c1-250
