cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
ok=0; bad=0
for i in $(seq 1 12); do
  if timeout 300 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_multi_device.py -m gpu -x -q -k "concurrent or dispatched or common" > gpurun_out/soak_$i.log 2>&1; then ok=$((ok+1)); else bad=$((bad+1)); tail -5 gpurun_out/soak_$i.log; fi
done
echo "soak: $ok ok, $bad failed"; tail -2 gpurun_out/soak_12.log
