# 1 -> 8 GPU scaling of bench.py (both arms at N=8), one box.  Usage: gpurun --gpus 8 -- bash scripts/gpu_scale_run.sh
cd $GRAFT_REPO_ROOT
T=${TAG:-r1k}
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
    bench.py --gpus $n --steps 5 --warmup 3 2>gpurun_out/scale_n$n.err | tail -1 > gpurun_out/${T}_bench_n$n.json
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
  bench.py --impl reference --gpus 8 --steps 3 --warmup 3 2>gpurun_out/scale_ref8.err | tail -1 > gpurun_out/${T}_bench_reference_n8.json
python - <<'PY'
import json
import os
T = os.environ.get("TAG", "r1k")
for f in (T + "_bench_n8", T + "_bench_n4", T + "_bench_n2", T + "_bench_reference_n8"):
    try:
        d = json.load(open("gpurun_out/" + f + ".json"))
        print(f, d["ms_per_step"], d["ms_per_view"], d["value"], d["e2e"]["value"], d["clocks"])
    except Exception as e:
        print(f, "failed", e)
PY
tail -3 gpurun_out/scale_n8.err
