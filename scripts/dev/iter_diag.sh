cd $GRAFT_REPO_ROOT
for v in "ITER_LAMBDA_NORMAL=0" "ITER_NO_REFL=1" "ITER_LAMBDA_NORMAL=0 ITER_NO_REFL=1"; do
  echo "== $v"
  env $v timeout 300 python tests/run_iteration.py c2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k: round(v,7) for k,v in d['first_iteration']['grad_rel_l2'].items()})"
done
