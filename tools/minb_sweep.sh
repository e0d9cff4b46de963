# launch-bound variants of tc_pipeline.cu (built by hand into csrc/variants/ with -DTC_*_MINB=n, see profiles/r1_minb_sweep.md)
TAG=base python tools/narrow_probe.py 4096 6
for f in trajopt-1_b200/csrc/variants/lib_*.so; do
  v=$(basename $f .so); TAG=${v#lib_} TC_LIB_PATH=$PWD/$f python tools/narrow_probe.py 4096 6
done
