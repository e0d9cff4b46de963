#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $out/r2_run9_pytest.log 2>&1
echo "pytest rc=$?"
grep -n "^FAILED\|^E  .*Error\|^E  .*assert\|passed\|failed" $out/r2_run9_pytest.log | head -30
: > $out/r2_run9_probe.log
TAG=spec timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run9_probe.log 2>&1
TC_NO_SPECULATION=1 TAG=exact timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run9_probe.log 2>&1
TC_PROBE_PREC=f32 TAG=f32 timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run9_probe.log 2>&1
TC_PROBE_EVAL=discrete TAG=lvs_discrete timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run9_probe.log 2>&1
grep tag= $out/r2_run9_probe.log
timeout 600 python bench.py --steps 10 --warmup 3 > $out/r2_run9_bench.json 2> $out/r2_run9_bench.err; echo "bench rc=$?"; tail -5 $out/r2_run9_bench.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_run9_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}); print(d['e2e']); print(d['stage_ms']); print(d['sqp_iteration_ms']); print(d['roofline']['frac'], d['roofline'].get('frac_executed')); print(json.dumps(d.get('other_configs'), indent=1)); print(d['cpu_baseline'])
PY
for c in 768 1536 2048; do
  timeout 300 python bench.py --steps 8 --warmup 3 --skip-extras --e2e-chunk $c > gpurun_out/r2_run9_e2e_c$c.json 2>/dev/null
  python -c "import json;d=json.load(open('gpurun_out/r2_run9_e2e_c$c.json'));print('chunk',$c,'e2e',d['e2e']['ms_per_step'],'dev',d['ms_per_step'])"
done
for wk in 2 4; do
  timeout 300 python bench.py --steps 8 --warmup 3 --skip-extras --e2e-workers $wk > gpurun_out/r2_run9_e2e_w$wk.json 2>/dev/null
  python -c "import json;d=json.load(open('gpurun_out/r2_run9_e2e_w$wk.json'));print('workers',$wk,'e2e',d['e2e']['ms_per_step'],'dev',d['ms_per_step'])"
done
