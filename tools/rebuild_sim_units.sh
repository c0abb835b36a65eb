#!/bin/bash
# experiment helper: rebuild the similarity kernel with a different B-ring depth and relink (run on the GPU box)
set -e
cd "$(dirname "$0")/.."
U=$1
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -cudart shared -DSIM_UNITS=$U -c onepose_b200/csrc/sim_dual_softmax.cu -o onepose_b200/lib/obj/sim_dual_softmax.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart shared -o onepose_b200/lib/libonepose_b200.so onepose_b200/lib/obj/*.o
