#!/bin/bash
# Developer harnesses (not part of the product). Binaries go to tools/_build/ (git-ignored, travel to the GPU box).
set -e
cd "$(dirname "$0")"
mkdir -p _build
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17"
nvcc $FLAGS -DDBA_CHOL_TRACE $EXTRA -o _build/chol_bench chol_bench.cu ../cut3r_slam_b200/csrc/dba_chol_fused.cu ../cut3r_slam_b200/csrc/dba_chol.cu ../cut3r_slam_b200/csrc/dba_trisolve_cluster.cu
nvcc $FLAGS -o _build/umma_test umma_test.cu
