set -x
cd $GRAFT_REPO_ROOT
timeout 600 python scripts/gpu_slices.py bench c4 c2 > gpurun_out/r2d_slices.jsonl 2> gpurun_out/r2d_slices.err; cat gpurun_out/r2d_slices.jsonl; tail -3 gpurun_out/r2d_slices.err
