#!/bin/bash
# Developer A/B builds: tools/build_variant.sh NAME [-DKNOB=VALUE ...]  ->  build/libogg_NAME.so  (select with OGG_B200_LIB)
NAME=$1; shift
mkdir -p build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -lineinfo -std=c++17 --shared -Xcompiler -fPIC -Xcompiler -O2 \
     -cudart shared -Xptxas -v "$@" -o build/libogg_$NAME.so ocean_model_grid_generator_b200/csrc/ogg_b200.cu 2>&1 \
  | grep -E "k_dp_hsamplesILi2|k_bp_quad_fusedILi5" -A3 | grep -E "spill|Used" | tr '\n' ' '
echo " <- $NAME"
