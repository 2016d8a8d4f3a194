# ncu --set full over every kernel family of the path, one warmed-up step each (scripts/ncu_target.py):
#   gpurun --timeout 2400 -- bash scripts/gpu_ncu_round.sh        (TAG=r2 by default)
# Only this library's kernels (namespace b200gs) are captured; the raw metric pages are exported as CSV on the box and
# the .ncu-rep files (tens of MB each) are dropped except the surfel one, so the results fit gpurun's 64 MiB return.
cd $GRAFT_REPO_ROOT
T=${TAG:-r2}
for tgt in ${TARGETS:-surfel gauss cubemap glue ssim adam}; do
  timeout 300 python scripts/ncu_target.py $tgt > /dev/null 2> gpurun_out/${T}_ncu_${tgt}.err || { echo "$tgt failed without ncu"; tail -3 gpurun_out/${T}_ncu_${tgt}.err; continue; }
  timeout 900 ncu --set full --clock-control none --profile-from-start off --kernel-name-base demangled -k regex:b200gs \
      -o /tmp/prof_${T}_${tgt} -f python scripts/ncu_target.py $tgt > gpurun_out/${T}_ncu_${tgt}.log 2>&1
  tail -1 gpurun_out/${T}_ncu_${tgt}.log
  ncu -i /tmp/prof_${T}_${tgt}.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_${tgt}_raw.csv 2>/dev/null
  ls -la /tmp/prof_${T}_${tgt}.ncu-rep
done
cp /tmp/prof_${T}_surfel.ncu-rep gpurun_out/ 2>/dev/null
# launch list of the bench command (shares of the step, cold-cache and serialised)
NCU_CMD="python bench.py --steps 2 --warmup 1 --views-per-step 1 --no-cpu-baseline --no-e2e"
timeout 300 $NCU_CMD > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_bench_steps2.csv $NCU_CMD > gpurun_out/${T}_ncu_l.log 2>&1
du -sh gpurun_out
