#!/bin/bash
# e2e leg of bench.py under pipeline tuning knobs (TC_PIPELINE_DEPTH, TC_PIPELINE_SCHEDULE, --e2e-chunk, --e2e-workers)
cd "$GRAFT_REPO_ROOT" || exit 1
run() { # label, workers, chunk, env...
  label=$1; w=$2; c=$3; shift 3
  env "$@" timeout 300 python bench.py --skip-extras --steps 10 --warmup 4 --e2e-workers $w --e2e-chunk $c 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', round(j['ms_per_step'],2), round(j['e2e']['ms_per_step'],2))"
}
run base 4 1024 X=1
run depth2 4 1024 TC_PIPELINE_DEPTH=2
run depth3 6 1024 TC_PIPELINE_DEPTH=3
run depth2_c768 6 768 TC_PIPELINE_DEPTH=2
run depth2_c1365 4 1365 TC_PIPELINE_DEPTH=2
run depth2_c2048 4 2048 TC_PIPELINE_DEPTH=2
run depth2_sched 6 1024 TC_PIPELINE_DEPTH=2 TC_PIPELINE_SCHEDULE=1536,1024,768,512,256
