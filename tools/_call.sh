cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
tools/bench_variants.sh "main vrow0" 24 C4 2>&1 | grep -v FATAL
