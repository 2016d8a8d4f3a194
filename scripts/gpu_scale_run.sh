# 1 -> N GPU scaling on one box:  gpurun --gpus 8 -- bash scripts/gpu_scale_run.sh     (NS="8 4 2" by default)
#   per N: the all-reduce check (own kernel vs ncclAllReduce), bench.py on the bench workload, bench.py on c5 (6M Gaussians)
cd $GRAFT_REPO_ROOT
T=${TAG:-r2}
for n in ${NS:-8 4 2}; do
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n"
  timeout 200 $RUN tests/run_dp_check.py 2>gpurun_out/${T}_dp_check_n$n.err | tail -1 > gpurun_out/${T}_dp_check_n$n.json
  timeout 300 $RUN bench.py --gpus $n --steps 5 --warmup 3 2>gpurun_out/${T}_scale_n$n.err | tail -1 > gpurun_out/${T}_bench_n$n.json
  timeout 300 $RUN bench.py --gpus $n --steps 5 --warmup 3 --nccl-allreduce --no-e2e 2>/dev/null | tail -1 > gpurun_out/${T}_bench_nccl_n$n.json
  [ -n "$SKIP_C5" ] || timeout 400 $RUN bench.py --gpus $n --steps 3 --warmup 3 --workload c5 --no-e2e 2>gpurun_out/${T}_scale_c5_n$n.err | tail -1 > gpurun_out/${T}_bench_c5_n$n.json
done
if [ -n "$WITH_REF" ]; then
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --impl reference --gpus 8 --steps 3 --warmup 3 2>gpurun_out/${T}_scale_ref8.err | tail -1 > gpurun_out/${T}_bench_reference_n8.json
fi
python - <<'PY'
import glob, json, os
T = os.environ.get("TAG", "r2")
for f in sorted(glob.glob(f"gpurun_out/{T}_bench*_n*.json")) + sorted(glob.glob(f"gpurun_out/{T}_dp_check_n*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        if "value" in d:
            sc = d.get("scale") or {}
            print(os.path.basename(f), "ms/step", d["ms_per_step"], "ms/view", d["ms_per_view"], "Mpix/s", d["value"], "e2e", (d.get("e2e") or {}).get("value"),
                  "| compute", sc.get("compute_ms_per_rank_min"), sc.get("compute_ms_per_rank_max"), "allreduce", sc.get("allreduce_ms"), sc.get("allreduce_algo", "")[:24])
        else:
            print(os.path.basename(f), {k: d[k] for k in ("algo", "own_kernel_ms", "nccl_ms", "p2p_ms", "ok") if k in d})
    except Exception as e:
        print(f, "failed", e)
PY
