cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
tests/bin/tma_row_probe > gpurun_out/r2w_tma_row_probe.txt 2>&1; cat gpurun_out/r2w_tma_row_probe.txt
( time timeout 900 python -m pytest tests/test_gpu_multi_device.py tests/test_link_compat.py tests/test_gpu_pygrackle.py -m gpu -x -q -rs ) > gpurun_out/r2w_pytest.log 2>&1; tail -5 gpurun_out/r2w_pytest.log
python tools/multi_device_probe.py 32 B > gpurun_out/r2w_multi_device.txt 2>&1
python tools/multi_device_probe.py 32 C >> gpurun_out/r2w_multi_device.txt 2>&1
cat gpurun_out/r2w_multi_device.txt | cut -c1-300
