#!/usr/bin/env bash
# Build the reference's own translation unit as a HOST program (parity oracle + timed CPU
# baseline).  TEST INFRASTRUCTURE ONLY.  Reads /root/reference (never copies its sources
# into the repository): the edited translation unit is a generated file under oracle/_ref/,
# which is git-ignored.  Edits made to the generated copy, all by line-addressed sed:
#   1. cu:244  the <<<grid,block>>> launch becomes SHIM_LAUNCH(...)  (every variant)
#   2. cu:577/589/592 and cu:621/633/648  erase-while-iterating (undefined behaviour, the
#      shipped -O2 build segfaults when the product holds an all-zero tile) becomes
#      collect-then-erase                                          ("fixed" variants only)
#   3. cu:162,179,224,241,256,273  the reference's own commented-out phase-timer prints are
#      re-enabled                                                  ("timed" variant only)
#   4. (ref_gpu only) no launch edit at all: the translation unit is compiled by nvcc for sm_100a with the
#      real CUDA runtime (edits 2+3 only), i.e. the reference's own kernel and host packing on the
#      same B200 -- SURVEY.md 8(d) "Reference GPU build on B200", a second reported comparator.
# Variants: ref_host (1), ref_host_fixed (1+2), ref_host_timed (1+2+3, OpenMP block loop off),
#           ref_host_timed_omp (1+2+3 and the emulated launch spread over OpenMP threads).
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
ref="${BSP_REFERENCE_DIR:-/root/reference}"
src="$ref/sparse_matrix_mult.cu"
out="$here/_ref"
want="b40ceddf0778746baa72a749ba9597c5da309bc20f6ccb9eb2fa1f52ec37fece"
if [ ! -f "$src" ]; then echo "build_ref: $src not present; keeping prebuilt oracle/_ref" >&2; exit 0; fi
have="$(sha256sum "$src" | cut -d' ' -f1)"
if [ "$have" != "$want" ]; then echo "build_ref: reference source changed (sha256 $have)" >&2; exit 1; fi
mkdir -p "$out"
launch='244s/matrix_multiplyKernel<<<grid,block>>>(/SHIM_LAUNCH(matrix_multiplyKernel,grid,block,/'
fix='577s/^/std::vector<std::pair<int,int>> dead_; /;589s/ans.erase(it.first);/dead_.push_back(it.first);/;592s/$/ for(auto \&dk_:dead_) ans.erase(dk_);/;621s/^/std::vector<std::pair<int,int>> dead_; /;633s/ans.erase(it.first);/dead_.push_back(it.first);/;648s/$/ for(auto \&dk_:dead_) ans.erase(dk_);/'
timers='162s@//@@;179s@//@@;224s@//@@;241s@//@@;256s@//@@;273s@//@@'
sed -e "$launch" "$src" > "$out/ref_host.cpp"
sed -e "$launch" -e "$fix" "$src" > "$out/ref_host_fixed.cpp"
sed -e "$launch" -e "$fix" -e "$timers" "$src" > "$out/ref_host_timed.cpp"
grep -q 'SHIM_LAUNCH(matrix_multiplyKernel' "$out/ref_host.cpp"
CXX="${BSP_CXX:-/usr/bin/g++}"
flags="-O2 -fopenmp -w -I$here/shim"
$CXX $flags "$out/ref_host.cpp" -o "$out/ref_host"
$CXX $flags "$out/ref_host_fixed.cpp" -o "$out/ref_host_fixed"
$CXX $flags "$out/ref_host_timed.cpp" -o "$out/ref_host_timed"
$CXX $flags -DSHIM_OMP_LAUNCH "$out/ref_host_timed.cpp" -o "$out/ref_host_timed_omp"
if command -v nvcc >/dev/null 2>&1; then
  # the shim directory also holds the cuda_runtime.h stand-in, which must NOT be seen by nvcc
  mkdir -p "$out/mpi_only" && cp "$here/shim/mpi.h" "$out/mpi_only/mpi.h"
  sed -e "$fix" -e "$timers" "$src" > "$out/ref_gpu.cu"
  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -I"$out/mpi_only" -Xcompiler -fopenmp -ccbin "$CXX" -w \
       "$out/ref_gpu.cu" -o "$out/ref_gpu"
  rm -rf "$out/ref_gpu.cu" "$out/mpi_only"
fi
rm -f "$out"/*.cpp
echo "build_ref: built $(ls "$out" | tr '\n' ' ')"
