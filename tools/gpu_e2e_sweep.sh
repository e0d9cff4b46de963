#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
run() { # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --skip-extras --steps 10 --warmup 4 --e2e-workers ${W:-4} 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', round(j['ms_per_step'],2), round(j['e2e']['ms_per_step'],2))"
}
run base X=1
run serial TC_PIPELINE_SERIAL=1
for sch in "2304,1152,512,128" "2048,1024,512,256,256" "1536,1024,768,512,256" "2304,1152,640" "1792,1280,768,256"; do
  run "serial:$sch" TC_PIPELINE_SERIAL=1 TC_PIPELINE_SCHEDULE=$sch
  run "conc:$sch" TC_PIPELINE_SCHEDULE=$sch
done
