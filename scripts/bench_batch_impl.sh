#!/bin/bash
# FP8 decode batch 1..64: tcgen05 GEMM with converter warps (PGK_TC_IMPL=1) vs raw E4M3 tiles by TMA into kind::f8f6f4 (2)
for impl in 1 2; do echo "== PGK_TC_IMPL=$impl PGK_TC_MIN_M=${PGK_TC_MIN_M:-5}"; PGK_TC_IMPL=$impl python scripts/bench_batch.py ${1:-8}; done
