cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for k in 0 2; do echo "== 2^24 key $k"; CONFIGS=C2,C3,C4 ORDER_KEY=$k python tools/bench_configs.py 24 B 2>&1 | grep -E "Mcell|rror"; done > gpurun_out/r2am_key.txt 2>&1
echo "== 2^27 C4 key 2" >> gpurun_out/r2am_key.txt; ORDER_KEY=2 CONFIGS=C4 REPS=2 python tools/bench_configs.py 27 B 2>&1 | grep -E "Mcell|rror" >> gpurun_out/r2am_key.txt
grep -v FATAL gpurun_out/r2am_key.txt
