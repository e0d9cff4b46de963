# launch-bound variants of tc_pipeline.cu (built by hand into csrc/variants/, see profiles/r1_minb_sweep.md)
TAG=base python tools/narrow_probe.py 4096 6
for v in fin6 fin7 fin8; do
  TAG=$v TC_LIB_PATH=$PWD/trajopt-1_b200/csrc/variants/lib_$v.so python tools/narrow_probe.py 4096 6
done
