set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_configs_gpu.py tests/test_surfel_gpu.py -x -q -m gpu -k "depth_sliced or baseline_config or surfel_gpu" > gpurun_out/dev_pytest_gpu.log 2>&1; tail -5 gpurun_out/dev_pytest_gpu.log
timeout 600 python scripts/gpu_slices.py bench > gpurun_out/dev_slices.jsonl 2> gpurun_out/dev_slices.err; cut -c1-900 gpurun_out/dev_slices.jsonl | head -3; tail -3 gpurun_out/dev_slices.err
python - <<'PY'
import sys, os
sys.path.insert(0, "tests")
import torch, harness as hz, scenes, b200gs
for name in ("bench", "c2"):
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    C = hz.new_surfel_C()
    b200gs.cull_stats(reset=True)
    f = hz.surfel_forward(C, sc, cam, exact=True, cull_audit=True)
    torch.cuda.synchronize()
    st = b200gs.cull_stats(reset=True)
    print(name, "audit stats: visited(hit) pairs", st[0], "with >=1 contributing lane", st[1], "lane contributions", st[2], "sum busiest lane per (warp,batch)", st[3], "(warp,batch)", st[4])
PY
