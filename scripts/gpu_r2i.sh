set -x
cd $GRAFT_REPO_ROOT
N=${NGPU:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $RUN tests/run_dp_check.py > gpurun_out/r2i_dp_check_n$N.json 2> gpurun_out/r2i_dp_check_n$N.err; tail -1 gpurun_out/r2i_dp_check_n$N.json; tail -3 gpurun_out/r2i_dp_check_n$N.err | cut -c1-300
