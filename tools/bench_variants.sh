#!/bin/bash
# Run tools/bench_configs.py against every library variant given (names of grackle_b200/lib_v*):
#   tools/bench_variants.sh "v0 vA vB" [log2 cells] [configs]
here=$(cd "$(dirname "$0")/.." && pwd)
for v in $1; do
  so="$here/grackle_b200/lib_$v/libgrackle_b200.so"
  [ "$v" = "main" ] && so="$here/grackle_b200/lib/libgrackle_b200.so"
  echo "== $v"
  GRB200_LIB="$so" CONFIGS="${3:-C4}" python "$here/tools/bench_configs.py" "${2:-24}" B 2>&1 | grep -E "Mcell|Error|error" 
done
