#!/bin/bash
# usage: tools/build_variant.sh NAME "EXTRA nvcc flags"
# Builds libgrt_sm100.so with extra compile-time flags in a scratch copy of the sources and
# puts it, with a copy of libgrtcode.so (which finds its sibling through rpath $ORIGIN), under
# grtcode_b200/variants/NAME/.  GRT_LIB_DIR=<that directory> makes api.py load it: A/B
# measurements of kernel variants without touching the shipped build.
set -e
NAME=$1; EXTRA=$2
ROOT=$(cd "$(dirname "$0")/.." && pwd)
W=/tmp/grt_variants/$NAME
rm -rf $W; mkdir -p $W/grtcode_b200 $ROOT/grtcode_b200/variants/$NAME
cp -r $ROOT/include $W/include
mkdir -p $W/grtcode_b200/csrc
cp $ROOT/grtcode_b200/csrc/*.cu $ROOT/grtcode_b200/csrc/*.cuh $ROOT/grtcode_b200/csrc/Makefile $W/grtcode_b200/csrc/
# only gas.cu depends on the sweep macros: reuse the other objects
cp $ROOT/grtcode_b200/csrc/pow5_table.inc $W/grtcode_b200/csrc/
cp $ROOT/grtcode_b200/csrc/context.o $ROOT/grtcode_b200/csrc/solvers.o $ROOT/grtcode_b200/csrc/peak.o $ROOT/grtcode_b200/csrc/hitran_parse.o $W/grtcode_b200/csrc/ 2>/dev/null || true
touch $W/grtcode_b200/csrc/context.o $W/grtcode_b200/csrc/solvers.o $W/grtcode_b200/csrc/peak.o $W/grtcode_b200/csrc/hitran_parse.o
make -C $W/grtcode_b200/csrc ../libgrt_sm100.so EXTRA="$EXTRA" > $W/build.log 2>&1 || (tail -20 $W/build.log; exit 1)
cp $W/grtcode_b200/libgrt_sm100.so $ROOT/grtcode_b200/variants/$NAME/
cp $ROOT/grtcode_b200/libgrtcode.so $ROOT/grtcode_b200/variants/$NAME/
cp $W/grtcode_b200/csrc/gas.ptxas.log $ROOT/grtcode_b200/variants/$NAME/gas.ptxas.log
echo "built variant $NAME ($EXTRA)"
