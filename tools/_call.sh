cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
tools/bench_variants.sh "vcl2 main" 24 C4 > gpurun_out/r2u_variants.txt 2>&1
tools/bench_variants.sh "vcl2 main" 24 C1,C2,C3 >> gpurun_out/r2u_variants.txt 2>&1
for v in lib_vcl2 lib; do echo "== $v" >> gpurun_out/r2u_variants.txt; GRB200_LIB=$PWD/grackle_b200/$v/libgrackle_b200.so python tools/calc_probe.py 16 >> gpurun_out/r2u_variants.txt 2>&1; done
cat gpurun_out/r2u_variants.txt
( time timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2u_pytest.log 2>&1; tail -4 gpurun_out/r2u_pytest.log
