# Round measurements on ONE B200:  gpurun --timeout 2400 -- bash scripts/gpu_measure_round.sh
# (tests, smoke, config table, both bench arms, update / loss harnesses, ncu launch list + full capture)
set -x
cd $GRAFT_REPO_ROOT
T=${TAG:-r1k}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 900 python tests/run_configs.py c1 c2 c3 c4 c3g c5 > gpurun_out/${T}_configs.jsonl 2> gpurun_out/${T}_configs.err
timeout 300 python tests/run_update.py > gpurun_out/${T}_update.jsonl 2>/dev/null
timeout 300 python tests/run_ssim.py > gpurun_out/${T}_ssim.jsonl 2>/dev/null
timeout 300 python tests/run_shade.py > gpurun_out/${T}_shade.jsonl 2>/dev/null
timeout 300 python tests/run_iteration.py > gpurun_out/${T}_iteration.jsonl 2>/dev/null
timeout 300 python tests/run_scene_io.py > gpurun_out/${T}_scene_io.jsonl 2>/dev/null
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 2>/dev/null | tail -1 > gpurun_out/${T}_bench_reference.json
timeout 600 python bench.py --steps 5 --warmup 3 2>/dev/null | tail -1 > gpurun_out/${T}_bench_new.json
timeout 300 python bench.py --steps 3 --warmup 3 --exact-lists --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/${T}_bench_new_exact.json
timeout 300 python bench.py --steps 20 --warmup 5 --views-per-step 1 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/${T}_bench_new_single_view.json
# ncu: one view per step keeps the launch list short; same kernels as the batch run
NCU_CMD="python bench.py --steps 2 --warmup 1 --views-per-step 1 --no-cpu-baseline --no-e2e"
timeout 300 $NCU_CMD > /dev/null 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_steps2.csv $NCU_CMD > gpurun_out/ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^(kept_rows|radix_|scan_|surfel_|tile_|depth_keys|duplicate)" -s 35 -c 21 -o gpurun_out/prof_${T} $NCU_CMD > gpurun_out/ncu_f.log 2>&1
tail -2 gpurun_out/ncu_f.log
cut -c1-300 gpurun_out/${T}_configs.jsonl
