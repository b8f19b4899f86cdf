#!/bin/bash
# Builds lib/libmossdec_<name>.so with one source recompiled under extra flags (A/B experiments; see gpu_ab.sh).
# usage: build_variant.sh <name> <file.cu> "<extra nvcc flags>"
set -e
cd "$(dirname "$0")/../moss_speech_decoder_cosy_b200/csrc"
NAME=$1; SRC=$2; EXTRA=$3
B=../../build/csrc
make -s -j16 all
mkdir -p $B/var
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I../../include $EXTRA -c $SRC -o $B/var/${NAME}.o
OBJS=$(ls $B/*.o | grep -v "/${SRC%.cu}.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/libmossdec_${NAME}.so $OBJS $B/var/${NAME}.o -lcudart
echo built ../lib/libmossdec_${NAME}.so
