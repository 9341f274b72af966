#!/bin/bash
# build an experimental variant of the library: scripts/build_variant.sh <name> <file.cu> <extra nvcc flags...>
# -> glue_factory_b200/build/variants/libb200glue_<name>.so (load with B200GLUE_LIB=...)
set -e
NAME=$1; SRC=$2; shift 2
cd "$(dirname "$0")/../glue_factory_b200"
mkdir -p build/variants
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --expt-extended-lambda --expt-relaxed-constexpr -Xcompiler -fPIC"
nvcc $FLAGS "$@" -c csrc/$SRC -o build/variants/${NAME}_${SRC%.cu}.o
OBJS=""
for o in build/*.o; do
  if [ "$(basename $o)" = "${SRC%.cu}.o" ]; then OBJS="$OBJS build/variants/${NAME}_${SRC%.cu}.o"; else OBJS="$OBJS $o"; fi
done
nvcc -shared -o build/variants/libb200glue_$NAME.so $OBJS -gencode arch=compute_100a,code=sm_100a
echo build/variants/libb200glue_$NAME.so
