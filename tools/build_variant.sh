#!/bin/bash
# tools/build_variant.sh NAME -DFOO=1 ...  ->  smc_physics_b200/libsmcphys_vNAME.so (kernel tuning experiments; not committed)
name=$1; shift
cd "$(dirname "$0")/../smc_physics_b200/csrc" && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -prec-sqrt=false -prec-div=false -ftz=true --shared -Xcompiler -fPIC "$@" -o ../libsmcphys_v$name.so smcphys.cu
