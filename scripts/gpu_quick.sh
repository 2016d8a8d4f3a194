cd $GRAFT_REPO_ROOT
timeout 600 python scripts/gpu_slices.py ${CFGS:-bench} 2> gpurun_out/quick.err | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if d['schedule'] in (None,):
        print(d['config'], d['fwd_bwd_ms'], d['fwd_ms'], {k: v[0] for k, v in d['stages'].items()})
"
tail -2 gpurun_out/quick.err
