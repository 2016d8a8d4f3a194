# Final-state validation on ONE B200:  gpurun --timeout 1200 -- bash scripts/gpu_validate.sh
# (GPU parity tests, smoke, both bench arms, whole-iteration and scene-IO harnesses; no profiler)
set -x
cd $GRAFT_REPO_ROOT
T=${TAG:-r1j}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 400 python bench.py --impl reference --steps 3 --warmup 3 2>/dev/null | tail -1 > gpurun_out/${T}_bench_reference.json
timeout 400 python bench.py --steps 5 --warmup 3 2>gpurun_out/${T}_bench_new.err | tail -1 > gpurun_out/${T}_bench_new.json
timeout 300 python tests/run_iteration.py > gpurun_out/${T}_iteration.jsonl 2>/dev/null
timeout 300 python tests/run_scene_io.py > gpurun_out/${T}_scene_io.jsonl 2>/dev/null
cut -c1-400 gpurun_out/${T}_bench_new.json gpurun_out/${T}_bench_reference.json
