#!/bin/bash
# build_variant.sh NAME SRC.cu FILE.cu [extra nvcc flags]  — link a variant library build/variants/lib_NAME.so in which SRC.cu
# (compiled with the extra flags, e.g. -DVCR_RF2_MINBLOCKS=8) replaces csrc/FILE.cu
set -e
cd "$(dirname "$0")/../vcr-gaus_b200"
mkdir -p build/variants
NAME=$1; SRC=$2; BASE=$(basename "$3" .cu); shift 3
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr -ccbin /usr/bin/g++ -Icsrc -I../include"
nvcc $FLAGS "$@" -c "$SRC" -o build/variants/$NAME.o 2> build/variants/$NAME.ptxas.log || { cat build/variants/$NAME.ptxas.log; exit 1; }
OBJS=$(ls build/*.o | grep -v "/$BASE.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/lib_$NAME.so $OBJS build/variants/$NAME.o -ccbin /usr/bin/g++
echo built build/variants/lib_$NAME.so
