#!/usr/bin/env bash
# Compiles the UNMODIFIED reference (alexander-sholohov/msk144cudecoder) from /root/reference/src,
# where the sources lie, into oracle/_ref/ (git-ignored, travels to the GPU box):
#   oracle/_ref/msk144cudecoder_ref   the reference CLI, its own kernels recompiled for sm_100a
#   oracle/_ref/ref_dump              oracle/ref_dump.cu + the reference's kernels (stage dumps)
# Flags follow the reference's CMake Release build (CMakeLists.txt:25-45,130-132): -O3 -DNDEBUG
# -DBIGSYM=1, no fast-math.  The one external symbol, gfortran's __packjt77_MOD_unpack77 (WSJT-X is
# an empty submodule), is supplied by msk144cudecoder_b200/csrc/host/unpack77.cpp.
# No reference source is copied; if /root/reference is absent (GPU box) this script does nothing.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${MSK144_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
HOST="$HERE/../msk144cudecoder_b200/csrc/host"
if [ ! -d "$REF/src" ]; then
  echo "build_ref: $REF/src not present; keeping prebuilt files in $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
NVCC="${NVCC:-nvcc}"
FLAGS=(-std=c++17 -O3 -DNDEBUG -DBIGSYM=1 -gencode arch=compute_100a,code=sm_100a --extended-lambda -I"$REF/src" -w)
COMMON=("$REF/src/analytic_fft.cu" "$REF/src/snr_tracker.cu" "$REF/src/decode_softbits.cpp" "$REF/src/metrics.cpp"
        "$REF/src/f_interop.cpp" "$REF/src/result_filter.cpp" "$HOST/unpack77.cpp")
if [ ! -x "$OUT/msk144cudecoder_ref" ] || [ "${FORCE:-0}" = 1 ]; then
  "$NVCC" "${FLAGS[@]}" -o "$OUT/msk144cudecoder_ref" "$REF/src/main.cu" "${COMMON[@]}" -lcufft
fi
if [ ! -x "$OUT/ref_dump" ] || [ "$HERE/ref_dump.cu" -nt "$OUT/ref_dump" ] || [ "${FORCE:-0}" = 1 ]; then
  "$NVCC" "${FLAGS[@]}" -o "$OUT/ref_dump" "$HERE/ref_dump.cu" "${COMMON[@]}" -lcufft
fi
ls -la "$OUT"
