#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
timeout 900 python -m pytest tests -m gpu -q > $out/r2_run11_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $out/r2_run11_pytest.log
: > $out/r2_run11_probe.log
TAG=tree timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run11_probe.log 2>&1
for v in t5 t4 t8 ev28 ev24 ev20; do
  TC_LIB_PATH=$PWD/scratch/variants/$v/csrc/libtrajopt_cuda.so TAG=$v timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run11_probe.log 2>&1
done
grep tag= $out/r2_run11_probe.log
