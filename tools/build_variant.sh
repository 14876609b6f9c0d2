#!/bin/bash
# build_variant.sh NAME FILE.cu   — link a variant library with build/variants/NAME.o replacing csrc/FILE.o
set -e
cd "$(dirname "$0")/../vcr-gaus_b200"
NAME=$1; SRC=$2; BASE=$(basename "$3" .cu)
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr -ccbin /usr/bin/g++ -Icsrc -I../include"
nvcc $FLAGS -c "$SRC" -o build/variants/$NAME.o 2> build/variants/$NAME.ptxas.log || { cat build/variants/$NAME.ptxas.log; exit 1; }
OBJS=$(ls build/*.o | grep -v "/$BASE.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/lib_$NAME.so $OBJS build/variants/$NAME.o -ccbin /usr/bin/g++
echo built build/variants/lib_$NAME.so
