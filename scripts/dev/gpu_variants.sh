# Development: time the ranking variants of the radix pass and the 3-CTA backward on ONE B200.
set -x
cd $GRAFT_REPO_ROOT
B="python bench.py --steps 20 --warmup 5 --views-per-step 1 --no-cpu-baseline --no-e2e"
for m in 0 1 2; do
  B200GS_RADIX_RANK=$m timeout 300 $B 2>/dev/null | tail -1 > gpurun_out/var_radix$m.json
done
B200GS_BWD_CTAS=3 timeout 300 $B 2>/dev/null | tail -1 > gpurun_out/var_bwd3.json
for m in 1 2; do
  B200GS_RADIX_RANK=$m timeout 600 python -m pytest tests/test_surfel_gpu.py tests/test_gauss_gpu.py -x -q -m gpu > gpurun_out/var_radix${m}_pytest.log 2>&1; tail -1 gpurun_out/var_radix${m}_pytest.log
done
B200GS_BWD_CTAS=3 timeout 600 python -m pytest tests/test_surfel_gpu.py -x -q -m gpu > gpurun_out/var_bwd3_pytest.log 2>&1; tail -1 gpurun_out/var_bwd3_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/var_*.json')):
    try:
        d=json.loads(open(f).read())
        st=d.get('stages',{})
        print(f, d['ms_per_step'], {k:round(v['ms'],4) for k,v in st.items()})
    except Exception as e: print(f, 'ERR', e)
PY
