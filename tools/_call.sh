cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
tools/bench_variants.sh "main vt640 vt384" 24 C4 > gpurun_out/r2c6_variants.log 2>&1; cat gpurun_out/r2c6_variants.log
