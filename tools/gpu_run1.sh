#!/bin/bash
# GPU session 1 of round 2: whole GPU suite on the restructured library, then stage times of the tuning builds
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $out/r2_run1_smi.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $out/r2_run1_pytest.log 2>&1
echo "pytest rc=$?" | tee -a $out/r2_run1_pytest.log
tail -5 $out/r2_run1_pytest.log
: > $out/r2_run1_probe.log
TAG=tree timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run1_probe.log 2>&1
for v in base0 gjk2 gjk4 run256 run1024 epa3 epa5 fkfma; do
  TC_LIB_PATH=$PWD/scratch/variants/$v/csrc/libtrajopt_cuda.so TAG=$v timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run1_probe.log 2>&1
done
cat $out/r2_run1_probe.log
