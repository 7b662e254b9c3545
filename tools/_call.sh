cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for v in lib lib_vda1; do
  echo "== $v HARNESS set H 2^21 / set B 2^23"
  GRB200_LIB=$PWD/grackle_b200/$v/libgrackle_b200.so CONFIGS=HARNESS python tools/bench_configs.py 21 H 2>&1 | grep -E "Mcell|rror"
  GRB200_LIB=$PWD/grackle_b200/$v/libgrackle_b200.so CONFIGS=HARNESS python tools/bench_configs.py 23 B 2>&1 | grep -E "Mcell|rror"
done > gpurun_out/r2ai_dustall.txt 2>&1
echo "== lib general C1,C2 (group size 1 now)" >> gpurun_out/r2ai_dustall.txt
CONFIGS=C1,C2 python tools/bench_configs.py 23 B general 2>&1 | grep -E "Mcell|rror" >> gpurun_out/r2ai_dustall.txt
cat gpurun_out/r2ai_dustall.txt
