#!/bin/bash
# Builds libccb.so for sm_100a (B200).  nvcc cross-compiles without a GPU.
# ccb_arm.cu and ccb_syn.cu are compiled once as the stateless translation unit and once per weight slot
# (-DCCB_SLOT=k, k = 0..7): every slot unit owns its __constant__ banks (see the top of ccb_arm.cu).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
OUT="$HERE/../_lib"
mkdir -p "$OUT" "$HERE/_obj"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
N_SLOTS=8
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I$ROOT/include -I$HERE ${CCB_EXTRA_FLAGS}"
JOBS=${CCB_BUILD_JOBS:-$(nproc)}
stale() {  # stale <object> <source>
  [ ! -f "$1" ] || [ "$2" -nt "$1" ] || [ "$HERE/ccb_internal.cuh" -nt "$1" ] || [ "$ROOT/include/ccb.h" -nt "$1" ] || [ "$HERE/build.sh" -nt "$1" ]
}
cmds=()
for f in ccb_api ccb_quant ccb_arm ccb_ups ccb_syn ccb_opt ccb_armint ccb_decint ccb_noalloc; do
  if stale "$HERE/_obj/$f.o" "$HERE/$f.cu"; then cmds+=("$NVCC $FLAGS -c $HERE/$f.cu -o $HERE/_obj/$f.o"); fi
done
for f in ccb_arm ccb_syn; do
  for ((k = 0; k < N_SLOTS; k++)); do
    if stale "$HERE/_obj/${f}_s$k.o" "$HERE/$f.cu"; then
      cmds+=("$NVCC $FLAGS -diag-suppress 177 -DCCB_SLOT=$k -c $HERE/$f.cu -o $HERE/_obj/${f}_s$k.o")
    fi
  done
done
if [ ${#cmds[@]} -gt 0 ]; then
  printf '%s\n' "${cmds[@]}" | xargs -P "$JOBS" -I{} bash -c '{}'
fi
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT/libccb.so" "$HERE"/_obj/*.o -lcudart
echo "built $OUT/libccb.so"
