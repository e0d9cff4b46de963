#!/bin/bash
# round-end evidence on one GPU: the GPU tests, the bench line (all extras), the reference arm, the smoke check,
# then the ncu launch list and one --set full capture (tools/gpu_profile.sh)
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
tag=${1:-r2_final}
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $out/${tag}_smoke.log
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_reference.json 2> $out/${tag}_reference.err; echo "reference rc=$?"
python - <<PY
import json
for f in ("$out/${tag}_bench.json", "$out/${tag}_reference.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k: j.get(k) for k in ("value", "ms_per_step", "gpu_launches")}, j.get("e2e"))
        if "roofline" in j: print(j["roofline"]); print(j.get("stage_ms")); print(j.get("sqp_iteration_ms")); print(j.get("fp32"))
    except Exception as e:
        print(f, "parse failed", e)
PY
bash tools/gpu_profile.sh
