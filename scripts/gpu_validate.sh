# Validation on ONE B200: all GPU tests, smoke, both bench arms (no profiler):  gpurun --timeout 2400 -- bash scripts/gpu_validate.sh
set -x
cd $GRAFT_REPO_ROOT
T=${TAG:-r2}
timeout 1500 python -m pytest tests -q -m gpu -s > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -40 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 500 python bench.py --steps 5 --warmup 3 2>gpurun_out/${T}_bench_new.err | tail -1 > gpurun_out/${T}_bench_new.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 3 2>gpurun_out/${T}_bench_reference.err | tail -1 > gpurun_out/${T}_bench_reference.json
cut -c1-600 gpurun_out/${T}_bench_new.json gpurun_out/${T}_bench_reference.json
tail -5 gpurun_out/${T}_bench_new.err
