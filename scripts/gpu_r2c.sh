set -x
cd $GRAFT_REPO_ROOT
T=${TAG:-r2c}
timeout 900 python -m pytest tests/test_configs_gpu.py tests/test_surfel_gpu.py -x -q -m gpu -k "depth_sliced or baseline_config or surfel_gpu" > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -25 gpurun_out/${T}_pytest_gpu.log
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>gpurun_out/${T}_bench_new.err | tail -1 > gpurun_out/${T}_bench_new.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c_bench_new.json'))
print(d["value"], d["ms_per_view"], d["e2e"]["ms_per_view"], d["forward_only"])
print({k:(v["ms"], v["launches_per_view"]) for k,v in d["stages"].items()})
PY
tail -5 gpurun_out/${T}_bench_new.err
