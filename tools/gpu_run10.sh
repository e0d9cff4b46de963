#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
: > $out/r2_run10_probe.log
for v in nosp sp3 sp4 sp5; do
  TC_LIB_PATH=$PWD/scratch/variants/$v/csrc/libtrajopt_cuda.so TAG=$v timeout 300 python tools/narrow_probe.py 4096 5 >> $out/r2_run10_probe.log 2>&1
done
grep tag= $out/r2_run10_probe.log
TC_LIB_PATH=$PWD/scratch/variants/sp4/csrc/libtrajopt_cuda.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x > $out/r2_run10_pytest.log 2>&1; echo "pytest(sp4) rc=$?"; tail -3 $out/r2_run10_pytest.log
