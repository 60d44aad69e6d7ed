#!/bin/bash
# Build an experiment copy of libssad_gen.so with extra -D flags for ssad_generate_pipeline.cu.
# usage: tools/build_variant.sh name -DSSAD_V7_UPT_FAST=16 ...   -> build/variants/libssad_<name>.so  (use with SSAD_LIB=)
set -e
cd "$(dirname "$0")/.."
P=self-supervised-anomaly-detection_b200
name=$1; shift
mkdir -p $P/build/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v "$@" \
     -c -o $P/build/variants/v7_$name.o $P/csrc/ssad_generate_pipeline.cu 2> $P/build/variants/v7_$name.ptxas
objs=$(ls $P/build/*.o | grep -v ssad_generate_pipeline.o)
nvcc -shared -o $P/build/variants/libssad_$name.so $objs $P/build/variants/v7_$name.o
echo $P/build/variants/libssad_$name.so
