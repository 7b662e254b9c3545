#!/bin/bash
# Build another copy of the CUDA library with extra compile flags (tuning experiments):
#   tools/build_variant.sh NAME "-DGRB_GROUP_WARPS_LEAN=1 ..."   ->  grackle_b200/lib_vNAME/libgrackle_b200.so
# Add -DGRB_DEV_ONLY_LEAN to compile only the compact device-resident solve instantiations.
set -e
here=$(cd "$(dirname "$0")/.." && pwd)
make -j8 -C "$here/grackle_b200/csrc" OUT="../lib_v$1" EXTRA="$2" 2>&1 | grep -i -E "error|warning: v" || true
ls -la "$here/grackle_b200/lib_v$1/libgrackle_b200.so"
