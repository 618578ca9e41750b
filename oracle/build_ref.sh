#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Compiles the REFERENCE's own hot-path CUDA kernels from where they lie
# under /root/reference (nothing is copied into this repo) with the reference's build flags
# (native/CMakeLists.txt:122: -O3 --use_fast_math --expt-relaxed-constexpr) plus the harness
# oracle/ref_harness.cu -> oracle/_ref/libref_cuda.so.  Skipped (exit 0) when the reference
# tree is not mounted (e.g. on the GPU box, which uses the prebuilt .so).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${PGK_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/native/ops" ]; then echo "build_ref: $REF not present, skipping"; exit 0; fi
mkdir -p "$OUT/obj"
OPS="$REF/native/ops"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 --use_fast_math --expt-relaxed-constexpr -std=c++17 -Xcompiler -fPIC -w)
SRCS=(
  "$OPS/matmul/gemv/w8a16_bf16/sm120/fp8_opt_kernels.cu"
  "$OPS/matmul/gemv/w4a16_bf16/sm120/nvf4_kernels.cu"
  "$OPS/matmul/gemv/w4a16_bf16/sm120/nvf4.cu"
  "$OPS/matmul/gemv/bf16_bf16/sm120/bf16_opt.cu"
  "$OPS/matmul/gemm/w8a16_bf16/sm120/w8a16_gemm.cu"
  "$OPS/matmul/gemm/w8a16_bf16/sm120/grouped_gemm.cu"
)
OBJS=()
for s in "${SRCS[@]}"; do
  o="$OUT/obj/$(basename "${s%.cu}").o"
  if [ ! -f "$o" ] || [ "$s" -nt "$o" ]; then nvcc "${FLAGS[@]}" -I"$OPS" -c "$s" -o "$o" & fi
  OBJS+=("$o")
done
h="$OUT/obj/ref_harness.o"
if [ ! -f "$h" ] || [ "$HERE/ref_harness.cu" -nt "$h" ]; then nvcc "${FLAGS[@]}" -I"$OPS" -c "$HERE/ref_harness.cu" -o "$h" & fi
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT/libref_cuda.so" "${OBJS[@]}" "$h"
echo "built $OUT/libref_cuda.so"
# The reference's pure-Python package, as it lies, for bench.py's config-1 CPU leg (GPT-2 small through the
# reference's own CausalTransformerModel.generate on its CPUSimulationBackend).  oracle/_ref/ is git-ignored
# (never part of this repo's history) but travels to the GPU box, where /root/reference does not exist.
if [ -d "$REF/src/pygpukit" ]; then
  rm -rf "$OUT/pysrc"; mkdir -p "$OUT/pysrc"
  cp -r "$REF/src/pygpukit" "$OUT/pysrc/pygpukit"
  find "$OUT/pysrc" -name '__pycache__' -type d -prune -exec rm -rf {} +
  echo "staged $OUT/pysrc/pygpukit"
fi
