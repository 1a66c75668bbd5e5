#!/bin/bash
# dev build: recompile only the translation units named (others keep their objects), then link.  tools/devbuild.sh wt_api wt_fused_flat
set -e
cd "$(dirname "$0")/../wavelettransforms_b200"
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -diag-suppress 177 -Xcompiler -fPIC $WT_NVCC_EXTRA"
pids=""
for f in "$@"; do $NV -c csrc/$f.cu -o build/$f.o & pids="$pids $!"; done
for p in $pids; do wait $p; done
/usr/local/cuda/bin/nvcc -shared -o lib/libwtprune.so build/*.o -gencode arch=compute_100a,code=sm_100a -lcudart
echo linked
