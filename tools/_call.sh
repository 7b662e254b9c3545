cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2s_pytest.log 2>&1; tail -4 gpurun_out/r2s_pytest.log
( time python bench.py > gpurun_out/r2s_bench_n1.json 2> gpurun_out/r2s_bench_n1.err ); tail -3 gpurun_out/r2s_bench_n1.err | cut -c1-300
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2s_bench_ref.json 2> gpurun_out/r2s_bench_ref.err )
B="--no-e2e --no-cpu --no-sweep --no-configs --no-set-c"
python bench.py --nk 16 --steps 2 --warmup 1 --no-cpu --no-configs --no-set-c > gpurun_out/r2s_nk16.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2s_launches.csv python bench.py --nk 16 --steps 2 --warmup 1 --no-cpu --no-configs --no-set-c > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve -c 1 -f -o gpurun_out/r2s_solve python bench.py --nk 16 --steps 1 --warmup 1 $B > gpurun_out/ncu2.log 2>&1; tail -2 gpurun_out/ncu2.log
python tools/calc_probe.py 16 > gpurun_out/r2s_calc_probe.txt 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_cooling_time -c 1 -f -o gpurun_out/r2s_ct python tools/calc_probe.py 16 cooling_time > gpurun_out/ncu3.log 2>&1; tail -2 gpurun_out/ncu3.log
cat gpurun_out/r2s_calc_probe.txt
ls -la gpurun_out
