cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q -rs ) > gpurun_out/r2ad_pytest.log 2>&1; tail -5 gpurun_out/r2ad_pytest.log
python tools/multi_device_probe.py 32 B > gpurun_out/r2ad_multi_device.txt 2>&1
python tools/multi_device_probe.py 32 C >> gpurun_out/r2ad_multi_device.txt 2>&1
cat gpurun_out/r2ad_multi_device.txt | cut -c1-300
