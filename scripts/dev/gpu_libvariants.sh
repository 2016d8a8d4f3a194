# Development: time alternative builds of the library (gaussian-splatting-reflection_b200/lib_var/libb200gs_<X>.so,
# compiled with other -DB200GS_*_CTAS values) on ONE B200.   LIBS="A B" bash scripts/dev/gpu_libvariants.sh
set -x
cd $GRAFT_REPO_ROOT
B="python bench.py --steps 20 --warmup 5 --views-per-step 1 --no-cpu-baseline --no-e2e"
timeout 300 $B 2>/dev/null | tail -1 > gpurun_out/var_default.json
timeout 300 python tests/run_configs.py c3g 2>/dev/null | tail -1 | cut -c1-300 > gpurun_out/var_default_c3g.json
for v in ${LIBS:-A}; do
  export B200GS_LIB=$GRAFT_REPO_ROOT/gaussian-splatting-reflection_b200/lib_var/libb200gs_$v.so
  timeout 300 $B 2>/dev/null | tail -1 > gpurun_out/var_$v.json
  timeout 300 python tests/run_configs.py c3g 2>/dev/null | tail -1 | cut -c1-300 > gpurun_out/var_${v}_c3g.json
  timeout 600 python -m pytest tests/test_surfel_gpu.py tests/test_gauss_gpu.py -x -q -m gpu > gpurun_out/var_${v}_pytest.log 2>&1; tail -1 gpurun_out/var_${v}_pytest.log
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/var_*.json')):
    try:
        d=json.loads(open(f).read())
        st=d.get('stages',{})
        print(f, d['ms_per_step'], {k:round(v['ms'],4) for k,v in st.items()})
    except Exception as e: print(f, open(f).read()[:300])
PY
