cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
tools/bench_variants.sh "main vsplit vt256 vfb4 vfb1 main" 24 C4 > gpurun_out/r2ab_variants.txt 2>&1
cat gpurun_out/r2ab_variants.txt
