#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Builds the UNMODIFIED reference VQ path (src/vq_encoder.cpp,
# src/bcn_compressor.cpp + vendored Compressonator static libs) from where it lies under
# $YUP_REFERENCE (default /root/reference) into oracle/_ref/libyupref.so, behind the C-ABI shim
# oracle/ref_harness.cpp.  Flags follow the reference's CMakeLists.txt:142 (-mavx2 -mfma, OpenMP);
# the reference's own CMake is not run (zstd submodule and TBB are missing).
# No reference source is copied; all outputs go to oracle/_ref/ (git-ignored, travels with gpurun).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${YUP_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$R/src" ]; then
  echo "build_ref: $R not present; keeping any prebuilt $OUT/libyupref.so" >&2
  exit 0
fi
mkdir -p "$OUT"
L="$R/thirdparty/Compressonator/linux/lib/x64"
CXXFLAGS="-std=c++17 -O2 -fopenmp -mavx2 -mfma -fPIC -I$R/include -I$R/thirdparty/colorm/include -I$R/thirdparty/Compressonator/include"
g++ $CXXFLAGS -c "$R/src/vq_encoder.cpp" -o "$OUT/vq_encoder.o"
g++ $CXXFLAGS -c "$R/src/bcn_compressor.cpp" -o "$OUT/bcn_compressor.o"
# src/vq_bcn_compressor.cpp (the container: Compress / DecompressToBCn) needs only <zstd.h>: a prototype-only header
# (oracle/zstd_shim) and the system libzstd.so.1 stand in for the missing submodule
ZLIB="$(ls /usr/lib/x86_64-linux-gnu/libzstd.so.1 /lib/x86_64-linux-gnu/libzstd.so.1 2>/dev/null | head -1 || true)"
EXTRA_OBJ=""
EXTRA_DEF=""
if [ -n "$ZLIB" ]; then
  g++ $CXXFLAGS -I"$HERE/zstd_shim" -c "$R/src/vq_bcn_compressor.cpp" -o "$OUT/vq_bcn_compressor.o"
  EXTRA_OBJ="$OUT/vq_bcn_compressor.o $ZLIB"
  EXTRA_DEF="-DYUPREF_HAVE_CONTAINER"
fi
g++ $CXXFLAGS $EXTRA_DEF -c "$HERE/ref_harness.cpp" -o "$OUT/ref_harness.o"
# -Bsymbolic: the reference's classes carry the same names as the drop-in's (libyuptexture.so); a process that loads both
# (tests/cpp/dropin_test.cpp) must see every call inside this library bind to the reference's own definitions
g++ -shared -fopenmp -Wl,-Bsymbolic -o "$OUT/libyupref.so" "$OUT/ref_harness.o" "$OUT/vq_encoder.o" "$OUT/bcn_compressor.o" $EXTRA_OBJ \
  -Wl,--start-group "$L/libCMP_Compressonator.a" "$L/libCMP_Common.a" "$L/libCMP_Core.a" \
  "$L/libCMP_Core_SSE.a" "$L/libCMP_Core_AVX.a" "$L/libCMP_Core_AVX512.a" -Wl,--end-group -lpthread -ldl
echo "build_ref: built $OUT/libyupref.so"
