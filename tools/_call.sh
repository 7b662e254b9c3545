cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time python bench.py --nk 8 --steps 1 --warmup 1 > gpurun_out/r2d3_bench_small.json 2> gpurun_out/r2d3_bench_small.err ); tail -5 gpurun_out/r2d3_bench_small.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2d3_bench_small.json"))
print("value %.3g e2e %.3g pageable %s" % (d["value"], d["e2e"]["value"], d["e2e_pageable"] and "%.3g" % d["e2e_pageable"]["value"]))
print("roofline", {k: d["roofline"][k] for k in ("achieved","peak","frac","peak_theoretical","frac_of_theoretical")})
print("sweep", {k: round(v["frac_hbm"],3) for k,v in d["calculate_sweep"].items()})
for k,v in d["configs"].items(): print(k, "%.3g cells/s" % v["cells_per_s"], "sub/cell %.1f" % v["subcycles_per_cell"], "frac %.3f" % v["roofline_frac_fp64"], "err", v.get("max_rel_err"), "itmax", v["cells_hit_itmax"])
print("set_c", d["set_c"]["contiguous_k_slabs"]["cells_per_s"], d["set_c"]["contiguous_k_slabs"]["subcycles_per_cell"])
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
PY
