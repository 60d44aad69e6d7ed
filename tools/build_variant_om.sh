#!/bin/bash
# Build an experiment copy of libssad_gen.so with extra -D flags for ssad_objmask.cu.
# usage: tools/build_variant_om.sh name -DSSAD_PLANES_WARPS=16 ...   -> build/variants/libssad_<name>.so  (use with SSAD_LIB=)
set -e
cd "$(dirname "$0")/.."
P=self-supervised-anomaly-detection_b200
name=$1; shift
mkdir -p $P/build/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v "$@" \
     -c -o $P/build/variants/om_$name.o $P/csrc/ssad_objmask.cu 2> $P/build/variants/om_$name.ptxas
objs=$(ls $P/build/*.o | grep -v ssad_objmask.o)
nvcc -shared -o $P/build/variants/libssad_$name.so $objs $P/build/variants/om_$name.o
echo $P/build/variants/libssad_$name.so
