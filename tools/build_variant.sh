#!/bin/bash
# A/B build of libfdgs.so: tools/build_variant.sh NAME FILE.cu "EXTRA NVCC FLAGS"  ->  tools/_bin/libfdgs_NAME.so
# (recompiles FILE.cu with the extra flags, links it with the current objects of the default build)
set -e
NAME=$1; FILE=$2; FLAGS=$3
PKG=4dgaussians-fast-train_b200
mkdir -p tools/_bin tools/_build
OBJ=tools/_build/${NAME}_${FILE%.cu}.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DFDGS_ARCH=100 $FLAGS \
  -c $PKG/csrc/$FILE -o $OBJ
OBJS=""
for o in $PKG/build/*.o; do
  if [ "$(basename $o)" == "${FILE%.cu}.o" ]; then OBJS="$OBJS $OBJ"; else OBJS="$OBJS $o"; fi
done
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o tools/_bin/libfdgs_${NAME}.so $OBJS -lcudart
echo tools/_bin/libfdgs_${NAME}.so
