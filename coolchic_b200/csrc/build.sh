#!/bin/bash
# Builds libccb.so for sm_100a (B200).  nvcc cross-compiles without a GPU.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
OUT="$HERE/../_lib"
mkdir -p "$OUT" "$HERE/_obj"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I$ROOT/include -I$HERE ${CCB_EXTRA_FLAGS}"
pids=()
for f in ccb_api ccb_quant ccb_arm ccb_ups ccb_syn ccb_opt ccb_armint ccb_decint; do
  if [ ! -f "$HERE/_obj/$f.o" ] || [ "$HERE/$f.cu" -nt "$HERE/_obj/$f.o" ] || [ "$HERE/ccb_internal.cuh" -nt "$HERE/_obj/$f.o" ] || [ "$ROOT/include/ccb.h" -nt "$HERE/_obj/$f.o" ]; then
    $NVCC $FLAGS -c "$HERE/$f.cu" -o "$HERE/_obj/$f.o" &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT/libccb.so" "$HERE"/_obj/*.o -lcudart
echo "built $OUT/libccb.so"
