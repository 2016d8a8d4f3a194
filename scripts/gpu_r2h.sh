# 2 GPUs: own all-reduce kernel vs NCCL, then the bench with either
set -x
cd $GRAFT_REPO_ROOT
T=${TAG:-r2h}
N=${NGPU:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $RUN tests/run_dp_check.py > gpurun_out/${T}_dp_check_n$N.json 2> gpurun_out/${T}_dp_check_n$N.err; tail -1 gpurun_out/${T}_dp_check_n$N.json; tail -5 gpurun_out/${T}_dp_check_n$N.err
timeout 400 $RUN bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err; tail -c 1500 gpurun_out/${T}_bench_n$N.json; tail -3 gpurun_out/${T}_bench_n$N.err
timeout 400 $RUN bench.py --gpus $N --steps 5 --warmup 3 --nccl-allreduce --no-e2e > gpurun_out/${T}_bench_nccl_n$N.json 2> gpurun_out/${T}_bench_nccl_n$N.err; tail -c 900 gpurun_out/${T}_bench_nccl_n$N.json
