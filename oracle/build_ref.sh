#!/bin/sh
# Compiles the reference's OWN source lines, where they lie under /root/reference, into
# oracle/_ref/ (git-ignored, travels to the GPU box).  No reference source is copied into
# the repo: the translation unit is assembled on a pipe and fed to the compiler's stdin.
#   src/tsdf.cu:15-60    GpuIntegrate body
#   src/tsdf.cu:170-218  SaveVoxelGrid2SurfacePointCloud   (TSDF:: qualifier stripped)
#   src/tsdf.cu:253-403  multiply_matrix / invert_matrix   (TSDF:: qualifier stripped)
set -e
REF=${REF_ROOT:-/root/reference}
SRC="$REF/src/tsdf.cu"
HERE=$(cd "$(dirname "$0")" && pwd)
OUT="$HERE/_ref"
mkdir -p "$OUT"
tu() {
  cat "$HERE/ref_prelude.inc"
  sed -n '15,60p' "$SRC"
  sed -n '170,218p;253,403p' "$SRC" | sed 's/TSDF:://'
  cat "$HERE/ref_epilogue.inc"
}
# strict: the parity pin for oracle/tsdf_oracle.c (same flags as liboracle.so)
tu | g++ -x c++ -std=c++11 -w -O2 -fPIC -shared -ffp-contract=off -o "$OUT/libtsdf_ref.so" -
# omp: the "reference CPU" timing arm (all host cores)
tu | g++ -x c++ -std=c++11 -w -O3 -march=x86-64-v3 -fopenmp -fPIC -shared -ffp-contract=off -o "$OUT/libtsdf_ref_omp.so" -
# gpu: the lifted kernel built by nvcc for sm_100a with nvcc's defaults (fmad on), as the
# reference's CMake would (CMakeLists.txt:18-21 has no flags at all)
if command -v nvcc >/dev/null 2>&1; then
  { sed -n '15,60p' "$SRC"; cat "$HERE/ref_gpu_epilogue.inc"; } > "$OUT/.tu.cu"
  nvcc -w -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared \
       -o "$OUT/libtsdf_ref_gpu.so" "$OUT/.tu.cu"
  rm -f "$OUT/.tu.cu"
fi
echo "oracle/_ref built: $(ls "$OUT")"
