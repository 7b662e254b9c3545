cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_gpu_variants.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
OV="cie_cooling=1,three_body_rate=4,h2_optical_depth_approximation=1,H2_self_shielding=3,self_shielding_method=3,h2_on_dust=1"
{ for c in C3 C4; do
echo "== $c + $OV : general forced"; CONFIGS=$c OVERRIDE=$OV python tools/bench_configs.py 23 B general 2>&1 | grep -E "Mcell|rror"
echo "== $c + same : default (VAR_COMMON)"; CONFIGS=$c OVERRIDE=$OV python tools/bench_configs.py 23 B 2>&1 | grep -E "Mcell|rror"
done
OV2="three_body_rate=1,cie_cooling=1"
echo "== C4 + $OV2 general / default"; CONFIGS=C4 OVERRIDE=$OV2 python tools/bench_configs.py 23 B general 2>&1 | grep -E "Mcell|rror"; CONFIGS=C4 OVERRIDE=$OV2 python tools/bench_configs.py 23 B 2>&1 | grep -E "Mcell|rror"
} > gpurun_out/r2an_common.txt 2>&1
grep -v FATAL gpurun_out/r2an_common.txt
