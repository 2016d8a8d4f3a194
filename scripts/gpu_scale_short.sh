# 8 / 4 / 2-GPU bench of this repo's arm only (short form of gpu_scale_run.sh).  gpurun --gpus 8 -- bash scripts/gpu_scale_short.sh
cd $GRAFT_REPO_ROOT
T=${TAG:-r1l}
for n in 8 4 2; do
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
    bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline 2>gpurun_out/scale_n$n.err | tail -1 > gpurun_out/${T}_bench_n$n.json
  cut -c1-330 gpurun_out/${T}_bench_n$n.json
done
