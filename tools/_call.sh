cd $GRAFT_REPO_ROOT
for v in v0 vc64m12 vc64m14 vc64m16 vc128m7 vc128m8; do
  so=$PWD/grackle_b200/lib_$v/libgrackle_b200.so; [ -f $so ] || continue
  echo "== $v"; GRB200_LIB=$so python tools/calc_probe.py 64 cooling_time 2>&1 | grep Gcell
done
echo "== solve u/s"; tools/bench_variants.sh "vu vs" 24 C4 2>&1 | grep -E "==|Mcell"
