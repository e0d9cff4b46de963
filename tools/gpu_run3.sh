#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $out/r2_run3_pytest.log 2>&1
echo "pytest rc=$?"
grep -n "^FAILED\|^E  .*Error\|passed\|failed" $out/r2_run3_pytest.log | head -30
TAG=spec timeout 300 python tools/narrow_probe.py 4096 5 > $out/r2_run3_probe.log 2>&1
TC_NO_SPECULATION=1 TAG=exact timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run3_probe.log 2>&1
cat $out/r2_run3_probe.log
timeout 600 python bench.py --steps 5 --warmup 3 > $out/r2_run3_bench.json 2> $out/r2_run3_bench.err; echo "bench rc=$?"; tail -c 3000 $out/r2_run3_bench.json
