set -x
timeout 300 python -m pytest tests/test_scene_io_gpu.py -x -q 2>&1 | tail -15
timeout 300 python tests/run_scene_io.py > gpurun_out/scene_io.jsonl 2> gpurun_out/scene_io.err; tail -3 gpurun_out/scene_io.err; cat gpurun_out/scene_io.jsonl
