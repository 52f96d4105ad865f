#!/bin/bash
# Builds libskp_b200.so in-tree for sm_100a (the .so travels to the GPU box with the snapshot).
set -e
cd "$(dirname "$0")/stablekeypoints_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared \
     -o ../libskp_b200.so *.cu "$@"
