#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
TC_NO_SPECULATION=1 TAG=dump TC_TRACE_DEEP=$PWD/$out/r2_deep_items.txt timeout 300 python tools/narrow_probe.py 512 1 > $out/r2_run5_probe.log 2>&1
tail -3 $out/r2_run5_probe.log; wc -l $out/r2_deep_items.txt
TAG=spec timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run5_probe.log 2>&1
TC_NO_SPECULATION=1 TAG=exact timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run5_probe.log 2>&1
grep tag= $out/r2_run5_probe.log
timeout 900 python -m pytest tests -m gpu -q -x > $out/r2_run5_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/r2_run5_pytest.log
