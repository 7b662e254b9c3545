cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for v in lib lib_vrare; do
  echo "== $v general"; GRB200_LIB=$PWD/grackle_b200/$v/libgrackle_b200.so python tools/bench_configs.py 23 B general 2>&1 | grep -E "Mcell|rror"
  echo "== $v default"; GRB200_LIB=$PWD/grackle_b200/$v/libgrackle_b200.so python tools/bench_configs.py 23 B 2>&1 | grep -E "Mcell|rror"
  GRB200_LIB=$PWD/grackle_b200/$v/libgrackle_b200.so CONFIGS=HARNESS python tools/bench_configs.py 21 H 2>&1 | grep -E "Mcell|rror"
done > gpurun_out/r2ak_rare.txt 2>&1
grep -v FATAL gpurun_out/r2ak_rare.txt
