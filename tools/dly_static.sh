#!/bin/bash
# Compile ONE instantiation of rhs_dly (isothermal MHD, reads the 2N register, no upwinding: the
# benchmarked substep-2/3 kernel) into a cubin in a few seconds and print registers/spills and the
# static issue-time model of its loops (tools/sass_stalls.py).  Used to rank code variants of the
# issue-bound kernel without GPU time.   usage: tools/dly_static.sh [extra nvcc flags]
here=$(cd "$(dirname "$0")" && pwd)
out=${TMPDIR:-/tmp}/dly_static
mkdir -p "$out"
inst=${DLY_INST:-"true, false, 0, true, false"}
cat > "$out/one.cu" <<EOT
#include "rhs_dly.cuh"
template __global__ void rhs_dly<$inst>(SubstepArgs, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap);
EOT
nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xptxas -v -I"$here/../pencil-code_b200/csrc" "$@" \
     -cubin -o "$out/one.cubin" "$out/one.cu" 2>&1 | grep -E "error|rhs_dly" -A2 | grep -E "error|registers|spill" | head -6
cd "$here" && python sass_stalls.py rhs_dly "$out/one.cubin" | cut -c1-330
