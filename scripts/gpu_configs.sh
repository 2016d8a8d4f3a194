cd $GRAFT_REPO_ROOT
T=${TAG:-r2l}
timeout 1200 python tests/run_configs.py ${CFGS:-c1 c2 c3 c4 c3g c5} > gpurun_out/${T}_configs.jsonl 2> gpurun_out/${T}_configs.err
python - <<PY
import json
for l in open("gpurun_out/${T}_configs.jsonl"):
    d = json.loads(l)
    print(d["config"], "new", round(d["new_ms"],3), "exact", round(d.get("new_exact_ms",0),3), "ref", round(d.get("ref_ms",0),3), "x", round(d.get("speedup",0),1), d.get("parity"))
PY
tail -3 gpurun_out/${T}_configs.err
