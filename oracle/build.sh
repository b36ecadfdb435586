#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds
#   oracle/_lib/libarmint_ref.so   the C restatement of the integer ARM (oracle/armint_ref.c)
# and, when the read-only reference checkout is present (this container; never the GPU box),
#   oracle/_ref/libccref.so        the reference's own decoder sources compiled IN PLACE + oracle/ref_driver.cpp
#   oracle/_ref/ccencapi*.so       the reference's pybind11 latent/weight entropy ENCODER, unmodified
# Nothing is copied out of /root/reference: the compiler reads the sources where they lie.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
mkdir -p "$HERE/_lib"
gcc -O2 -std=c11 -fPIC -shared -Wall -o "$HERE/_lib/libarmint_ref.so" "$HERE/armint_ref.c"

REF="${CCB_REFERENCE_ROOT:-/root/reference}/coolchic/cpp"
if [ -d "$REF" ]; then
  mkdir -p "$HERE/_ref"
  CXXFLAGS="-O2 -std=c++17 -fPIC -w -include cstdint -DCCDECAPI_CPU -I$REF"
  # decoder side: bitstream parsing, CABAC, contexts, frame decoder (ARM / upsampling / synthesis, scalar build)
  SRCS="cc-bitstream.cpp cc-contexts.cpp cc-frame-decoder.cpp frame-memory.cpp arm_cpu.cpp ups_cpu.cpp syn_cpu.cpp \
        BitStream.cpp TDecBinCoderCABAC.cpp Contexts.cpp"
  if [ ! -f "$HERE/_ref/libccref.so" ] || [ "$HERE/ref_driver.cpp" -nt "$HERE/_ref/libccref.so" ]; then
    ( cd "$REF" && g++ $CXXFLAGS -shared -o "$HERE/_ref/libccref.so" "$HERE/ref_driver.cpp" $SRCS )
  fi
  PYINC="$(python - <<'PY'
import sysconfig, pybind11
print("-I" + sysconfig.get_paths()["include"], "-I" + pybind11.get_include())
PY
)"
  EXT="$(python -c 'import sysconfig; print(sysconfig.get_config_var("EXT_SUFFIX"))')"
  if [ ! -f "$HERE/_ref/ccencapi$EXT" ]; then
    ( cd "$REF" && g++ $CXXFLAGS $PYINC -shared -o "$HERE/_ref/ccencapi$EXT" ccencapi.cpp cc-contexts.cpp BitStream.cpp \
        TEncBinCoderCABAC.cpp TDecBinCoderCABAC.cpp Contexts.cpp )
  fi
fi
echo "oracle build ok"
