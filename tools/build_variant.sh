#!/bin/bash
# Builds a variant of the library into tools/bin (git-ignored): tools/build_variant.sh <name> <extra nvcc flags...>
set -e
REPO=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; shift
OBJ=$REPO/build/variant_$NAME
mkdir -p $OBJ $REPO/tools/bin
for f in $REPO/rpvg_b200/csrc/*.cu; do
  b=$(basename $f .cu)
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I$REPO/include -ccbin /usr/bin/g++ "$@" -c $f -o $OBJ/$b.o &
done
wait
/usr/local/cuda/bin/nvcc -shared -o $REPO/tools/bin/librpvg_b200_$NAME.so $OBJ/*.o
echo built tools/bin/librpvg_b200_$NAME.so
