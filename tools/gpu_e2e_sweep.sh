#!/bin/bash
# e2e leg of bench.py under pipeline tuning knobs (TC_PIPELINE_DEPTH, TC_PIPELINE_SCHEDULE, --e2e-chunk, --e2e-workers)
cd "$GRAFT_REPO_ROOT" || exit 1
run() { # label, workers, chunk, env...
  label=$1; w=$2; c=$3; shift 3
  env "$@" timeout 300 python bench.py --skip-extras --steps 10 --warmup 4 --e2e-workers $w --e2e-chunk $c 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', round(j['ms_per_step'],2), round(j['e2e']['ms_per_step'],2))"
}
run base 4 1024 X=1
for sch in "1024,1024,1024,1024" "1024,1024,1024,768,256" "1024,1024,1024,512,512" "1280,1280,1024,512" "1024,1024,1024,512,384,128" "1152,1152,1024,512,256"; do
  run "sched:$sch" 4 1024 TC_PIPELINE_SCHEDULE=$sch
done
