#!/bin/bash
# Developer experiment: libdba with csrc/dba_linearize_v2.cu compiled under extra -D flags.
#   tools/build_v2_variant.sh <name> <flags...>   -> tools/_build/libdba_<name>.so
# Needs the product's objects (python -m cut3r_slam_b200.build).  On the GPU box:
#   DBA_LIB_PATH=tools/_build/libdba_<name>.so python bench.py --workload backend --brief
set -e
cd "$(dirname "$0")"
mkdir -p _build
name=$1; shift
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O3"
OBJ=../cut3r_slam_b200/build
OTHERS=$(ls $OBJ/*.o | grep -v dba_linearize_v2.o)
nvcc $FLAGS "$@" -I../cut3r_slam_b200/csrc -x cu -c ${V2_SRC:-../cut3r_slam_b200/csrc/dba_linearize_v2.cu} -o _build/v2_$name.o
nvcc -shared -o _build/libdba_$name.so $OTHERS _build/v2_$name.o -gencode arch=compute_100a,code=sm_100a -ldl
rm -f _build/v2_$name.o
