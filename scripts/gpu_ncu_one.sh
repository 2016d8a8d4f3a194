# ncu --set full of selected kernels of one single-view bench step:  KREGEX=... SKIP=.. COUNT=.. TAG=.. bash scripts/gpu_ncu_one.sh
cd $GRAFT_REPO_ROOT
NCU_CMD="python bench.py --steps 1 --warmup 1 --views-per-step 1 --no-cpu-baseline --no-e2e ${BENCH_ARGS:-}"
timeout 300 $NCU_CMD > /dev/null 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX}" -s ${SKIP:-0} -c ${COUNT:-4} -o gpurun_out/prof_${TAG} $NCU_CMD > gpurun_out/ncu_${TAG}.log 2>&1
tail -2 gpurun_out/ncu_${TAG}.log
