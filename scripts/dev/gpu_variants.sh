# Development: time variants selected by environment switches on ONE B200.
#   VARS="NAME=VALUE ..." bash scripts/dev/gpu_variants.sh      (each entry: one bench run + the surfel/3DGS parity tests)
set -x
cd $GRAFT_REPO_ROOT
B="python bench.py --steps 20 --warmup 5 --views-per-step 1 --no-cpu-baseline --no-e2e"
for v in ${VARS:-B200GS_NONE=0}; do
  env $v timeout 300 $B 2>/dev/null | tail -1 > gpurun_out/var_$v.json
  env $v timeout 600 python -m pytest tests/test_surfel_gpu.py tests/test_gauss_gpu.py -x -q -m gpu > gpurun_out/var_${v}_pytest.log 2>&1; tail -1 gpurun_out/var_${v}_pytest.log
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/var_*.json')):
    try:
        d=json.loads(open(f).read())
        st=d.get('stages',{})
        print(f, d['ms_per_step'], {k:round(v['ms'],4) for k,v in st.items()})
    except Exception as e: print(f, 'ERR', e)
PY
