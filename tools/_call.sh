cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi_device.py -m gpu -x -q 2>&1 | tail -3
python tools/multi_device_probe.py 32 B > gpurun_out/r2z_multi_device.txt 2>&1
python tools/multi_device_probe.py 32 C >> gpurun_out/r2z_multi_device.txt 2>&1
cat gpurun_out/r2z_multi_device.txt | cut -c1-300
