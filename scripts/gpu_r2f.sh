set -x
cd $GRAFT_REPO_ROOT
NCU_CMD="python bench.py --steps 1 --warmup 1 --views-per-step 1 --no-cpu-baseline --no-e2e"
timeout 300 $NCU_CMD > /dev/null 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"^(kept_rows|surfel_composite)" -s 14 -c 8 -o gpurun_out/prof_r2f $NCU_CMD > gpurun_out/ncu_r2f.log 2>&1
tail -3 gpurun_out/ncu_r2f.log
ls -la gpurun_out/prof_r2f.ncu-rep
