cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2aa_smoke.txt 2>&1; tail -3 gpurun_out/r2aa_smoke.txt | cut -c1-300
echo "== general instantiation forced, Set B, 16 Mi cells" > gpurun_out/r2aa_general.txt
python tools/bench_configs.py 24 B general 2>&1 | grep -E "Mcell|rror" >> gpurun_out/r2aa_general.txt
echo "== default (compact) instantiations" >> gpurun_out/r2aa_general.txt
python tools/bench_configs.py 24 B 2>&1 | grep -E "Mcell|rror" >> gpurun_out/r2aa_general.txt
cat gpurun_out/r2aa_general.txt
