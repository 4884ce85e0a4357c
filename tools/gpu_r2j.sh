#!/bin/bash
# round 2, pass j: ncu --set full of the SIREN FP32-FMA kernel with the branch-free sincos at the published 3x64 network,
# saturating grid (64 tasks x 16384 + 16384 points): compare with profiles/r01_ncu_full_pde_kernel_3x64_* (library sincosf)
export MPDE_W=64 MPDE_L=3 MPDE_T=64 MPDE_N=16384 MPDE_ACT=sin
bash tools/gpu_ncu_full.sh r02_j_siren grad pde_kernel loss pde_kernel
