set -x
cd $GRAFT_REPO_ROOT
timeout 900 python tests/run_configs.py c1 c2 c3 c4 c3g c5 > gpurun_out/r1g_configs.jsonl 2> gpurun_out/r1g_configs.err
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/r1g_bench_reference.json
timeout 300 python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/r1g_bench_new.json
timeout 300 python bench.py --steps 20 --warmup 5 --exact-lists --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/r1g_bench_new_exact.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches_bench_steps2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^(kept_rows|radix_|scan_|surfel_|tile_|depth_keys)" -s 21 -c 21 -o gpurun_out/prof_r1g python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_f.log 2>&1
tail -2 gpurun_out/ncu_f.log
cat gpurun_out/r1g_configs.jsonl | cut -c1-400
