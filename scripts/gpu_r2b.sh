set -x
cd $GRAFT_REPO_ROOT
T=${TAG:-r2b}
timeout 900 python -m pytest tests/test_render_gpu.py tests/test_configs_gpu.py -q -m gpu -s -k "render or mask or surgery or glue" > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -30 gpurun_out/${T}_pytest_gpu.log
timeout 600 python scripts/gpu_list_usage.py bench c3 c4 c2 c5 > gpurun_out/${T}_list_usage.jsonl 2> gpurun_out/${T}_list_usage.err; cat gpurun_out/${T}_list_usage.jsonl; tail -3 gpurun_out/${T}_list_usage.err
