#!/bin/bash
# build_variant.sh <name> [nvcc -D flags...] — librbacknn_<name>.so under gpurun_tmp/ with extra defines (kernel experiments;
# select with RBACKNN_LIB=gpurun_tmp/librbacknn_<name>.so).  Only rk_scan.cu is rebuilt unless FILES is set.
set -e
cd "$(dirname "$0")/.."
N=$1; shift
mkdir -p gpurun_tmp/obj_$N
P=vectorsearch-rbac_b200
FL="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr"
OBJS=""
for f in rk_api rk_search rk_scan rk_merge rk_util rk_ivf rk_ivf_kernels rk_bscan rk_bscan2 rk_tensor rk_exchange rk_multi rk_noalloc rk_assign_tc rk_pgpage; do
  if echo " ${FILES:-rk_scan} " | grep -q " $f "; then
    /usr/local/cuda/bin/nvcc $FL "$@" -c $P/csrc/$f.cu -o gpurun_tmp/obj_$N/$f.o &
    OBJS="$OBJS gpurun_tmp/obj_$N/$f.o"
  else
    OBJS="$OBJS $P/build/$f.o"
  fi
done
wait
/usr/local/cuda/bin/nvcc -shared -o gpurun_tmp/librbacknn_$N.so $OBJS -gencode arch=compute_100a,code=sm_100a -lcudart
echo gpurun_tmp/librbacknn_$N.so
