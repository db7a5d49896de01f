#!/bin/bash
# Build an experimental variant of libb200lab.so: tools/variant.sh NAME unit.cu "-DFOO=1 ..." -> tools/_bin/lib_NAME.so
# (the other units are taken from the regular in-tree objects; run `python -m isaaclab_b200.build` first).
# On the GPU box: cp tools/_bin/lib_NAME.so isaaclab_b200/libb200lab.so before the bench run.
set -e
NAME=$1; UNIT=$2; DEFS=$3
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OBJ=$ROOT/isaaclab_b200/csrc/_obj
mkdir -p $ROOT/tools/_bin
EXTRA=""
case $UNIT in gae.cu|mdp.cu|raycast.cu) EXTRA="-fmad=false";; esac
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -O2 --expt-relaxed-constexpr \
  $EXTRA $DEFS -c $ROOT/isaaclab_b200/csrc/$UNIT -o $ROOT/tools/_bin/${NAME}_${UNIT%.cu}.o
OBJS=""
for o in $OBJ/*.o; do
  if [ "$(basename $o)" == "${UNIT%.cu}.o" ]; then OBJS="$OBJS $ROOT/tools/_bin/${NAME}_${UNIT%.cu}.o"; else OBJS="$OBJS $o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $ROOT/tools/_bin/lib_$NAME.so $OBJS
echo $ROOT/tools/_bin/lib_$NAME.so
