#!/usr/bin/env bash
# Builds libp2p_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr
       -Xptxas -v)
mkdir -p obj
pids=()
for f in api conv_simt conv_tc conv_halo conv_wgrad_tc conv_wgrad_halo conv_co1 elementwise losses_optim weights; do
  if [ ! -f obj/$f.o ] || [ $f.cu -nt obj/$f.o ] || [ common.cuh -nt obj/$f.o ] || [ tc_common.cuh -nt obj/$f.o ] \
     || [ ../../include/p2p_b200.h -nt obj/$f.o ]; then
    ( "$NVCC" "${FLAGS[@]}" -c $f.cu -o obj/$f.o > obj/$f.log 2>&1 || { cat obj/$f.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
"$NVCC" -shared -o libp2p_b200.so obj/*.o -lcudart
echo "built $(pwd)/libp2p_b200.so"
