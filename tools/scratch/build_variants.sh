#!/bin/bash
# Builds variant libraries HERE (nvcc cross-compiles), so the GPU box only runs them:
#   tools/scratch/build_variants.sh name1 "-DFLAGS" name2 "-DFLAGS" ...   ->  tools/scratch/variants/<name>.so
# and on the box:  tools/scratch/run_variants.sh name1 name2 ...
set -e
cd "$(dirname "$0")/../../katib_b200/csrc"
make -j8 > /dev/null
mkdir -p ../../tools/scratch/variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC $flags -c kb_tpe_score.cu -o /tmp/kbv_$name.o
  objs=$(ls build/*.o | grep -v kb_tpe_score.o)
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tools/scratch/variants/$name.so $objs /tmp/kbv_$name.o -ldl
  echo built $name "($flags)"
done
