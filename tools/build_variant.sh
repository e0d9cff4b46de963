#!/bin/bash
# tools/build_variant.sh NAME "<nvcc -D flags>" [STRICT_FLAGS]  — tuning builds of libtrajopt_cuda.so under scratch/variants/NAME
# (selected at run time with TC_LIB_PATH; scratch/ is git-ignored but travels to the GPU box)
set -e
root=$(cd "$(dirname "$0")/.." && pwd)
dst=$root/scratch/variants/$1
mkdir -p "$dst/csrc" "$dst/include"
cp "$root"/trajopt-1_b200/csrc/*.cu "$root"/trajopt-1_b200/csrc/*.cuh "$root"/trajopt-1_b200/csrc/*.hpp "$root"/trajopt-1_b200/csrc/Makefile "$dst/csrc/"
sed -i 's|../../include/trajopt_cuda.h|'"$root"'/include/trajopt_cuda.h|g' "$dst"/csrc/*.hpp "$dst"/csrc/*.cu "$dst"/csrc/*.cuh "$dst"/csrc/Makefile
if [ $# -ge 3 ]; then
  make -C "$dst/csrc" -j4 -s VARIANT="$2" STRICT_FLAGS="$3"
else
  make -C "$dst/csrc" -j4 -s VARIANT="$2"
fi
ls -la "$dst/csrc/libtrajopt_cuda.so"
