#!/bin/sh
# Compiles the UNMODIFIED reference CUDA sources of the skeleton path (where they lie under
# /root/reference) into oracle/_ref/libcusk_ref.so for sm_100, to time the reference's own
# cuPC-S kernels on the same B200 beside ours.  Checker / baseline only:
# the product never links or loads it.  Outputs go to oracle/_ref/ (git-ignored, travels to the
# GPU box).  The reference's build system (cmake + Boost + googletest) is not used.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference/cusk
OUT="$HERE/_ref"
[ -d "$REF" ] || { echo "build_ref: $REF not present, skipping"; exit 0; }
mkdir -p "$OUT"
if [ -f "$OUT/libcusk_ref.so" ] && [ "$OUT/libcusk_ref.so" -nt "$REF/src/cuPC-S.cu" ] && [ -f "$OUT/libcusk_ref_corr.so" ]; then exit 0; fi
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-std=c++17 -arch=sm_100 -O3 --use_fast_math --expt-relaxed-constexpr -Xcompiler -fPIC -I$REF/include -w"
$NVCC $FLAGS -dc "$REF/src/cuPC-S.cu" -o "$OUT/cuPC-S.o" &
$NVCC $FLAGS -dc "$REF/src/hetcor-cuPC-S.cu" -o "$OUT/hetcor-cuPC-S.o" &
wait
$NVCC -arch=sm_100 -shared -o "$OUT/libcusk_ref.so" "$OUT/cuPC-S.o" "$OUT/hetcor-cuPC-S.o" -lcudart
# the reference's correlation kernels and their host drivers (src/corr_kernels.cu, src/corr_host.cu: C++ linkage,
# called through their mangled names by tests/ref_lib.py), same flags as its CMake build except the architecture
FLAGS20="-std=c++20 -arch=sm_100 -O3 --use_fast_math --expt-relaxed-constexpr -Xcompiler -fPIC -I$REF/include -w"
$NVCC $FLAGS20 -dc "$REF/src/corr_kernels.cu" -o "$OUT/corr_kernels.o" &
$NVCC $FLAGS20 -dc "$REF/src/corr_host.cu" -o "$OUT/corr_host.o" &
wait
$NVCC -arch=sm_100 -shared -o "$OUT/libcusk_ref_corr.so" "$OUT/corr_kernels.o" "$OUT/corr_host.o" -lcudart
rm -f "$OUT"/*.o
echo "build_ref: $OUT/libcusk_ref.so $OUT/libcusk_ref_corr.so"
