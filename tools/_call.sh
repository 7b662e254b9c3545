cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2g_bench_n2.json 2> gpurun_out/r2g_bench_n2.err ); tail -3 gpurun_out/r2g_bench_n2.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2g_bench_n2.json"))
print({k:d[k] for k in ("value","n_gpus","ms_per_step","scaling","gpu_launches")}); print("e2e", d["e2e"]["value"]); print("imbalance", d.get("imbalance",{}).get("max_over_mean")); sc=d["set_c"]; print("set_c", sc["contiguous_k_slabs"]["imbalance_max_over_mean"], sc["contiguous_k_slabs"]["cells_per_s"], sc["interleaved"] and (sc["interleaved"]["imbalance_max_over_mean"], sc["interleaved"]["cells_per_s"]))
PY
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2g_bench_ref_n2.json 2>/dev/null ); cut -c1-200 gpurun_out/r2g_bench_ref_n2.json
