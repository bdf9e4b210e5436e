#!/bin/bash
# profiles/build_variant.sh NAME [-Dmacro=val ...] : builds build/variants/NAME.so for A/B timing
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -Iinclude "$@" \
  -o build/variants/$name.so aircraft_b200/csrc/airc_api.cu -Xptxas -v 2>&1 | grep -A2 "${VKERNEL:-rollout_kernelILb1ELb0}" | grep -E "registers|spill" | tr '\n' ' '
echo " <- $name"
