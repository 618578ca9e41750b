// moe.cu -- MoE routing and expert GEMV for decode (num_tokens <= 64).
// replaces native/ops/moe/{topk,permute,moe}_kernels.cuh as bound by native/bindings/moe.cpp:43-223
// and grouped_gemm_tiled_kernel (native/ops/matmul/gemm/w8a16_bf16/sm120/grouped_gemm.cu:77-201).
// Index work is bit-exact; the permutation is built in ascending flat-index order (one of the
// orders the reference's atomics can produce; everything downstream is invariant to it).
#include "common.cuh"
#include "gemv_common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

template <int DT> __device__ __forceinline__ float mld(const void* p, size_t i);
template <> __device__ __forceinline__ float mld<0>(const void* p, size_t i) { return ((const float*)p)[i]; }
template <> __device__ __forceinline__ float mld<1>(const void* p, size_t i) { return __half2float(((const __half*)p)[i]); }
template <> __device__ __forceinline__ float mld<2>(const void* p, size_t i) { return bf16_bits_to_f32(((const uint16_t*)p)[i]); }
template <int DT> __device__ __forceinline__ void mst(void* p, size_t i, float v);
template <> __device__ __forceinline__ void mst<0>(void* p, size_t i, float v) { ((float*)p)[i] = v; }
template <> __device__ __forceinline__ void mst<1>(void* p, size_t i, float v) { ((__half*)p)[i] = __float2half_rn(v); }
template <> __device__ __forceinline__ void mst<2>(void* p, size_t i, float v) { ((uint16_t*)p)[i] = f32_to_bf16_bits(v); }

// topk_with_indices_*_kernel topk_kernels.cuh:69-186: k passes of strict '>' argmax (lowest
// index wins ties).  One warp per token: lane-strided scan + (value, index) shuffle reduce.
template <int DT>
__global__ void moe_topk_kernel(const void* __restrict__ logits, void* __restrict__ values,
                                int32_t* __restrict__ indices, int T, int E, int k) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31;
    const int t = (int)(((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (t >= T) return;
    constexpr int MAXPL = 16;   // up to 512 experts
    float loc[MAXPL];
#pragma unroll
    for (int i = 0; i < MAXPL; ++i) { int e = lane + 32 * i; loc[i] = (e < E) ? mld<DT>(logits, (size_t)t * E + e) : -INFINITY; }
    for (int j = 0; j < k; ++j) {
        float bv = -1e30f;
        int bi = 0x7fffffff;
#pragma unroll
        for (int i = 0; i < MAXPL; ++i) { int e = lane + 32 * i; if (e < E && loc[i] > bv) { bv = loc[i]; bi = e; } }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (bi == 0x7fffffff) bi = 0;
        if (lane == 0) { mst<DT>(values, (size_t)t * k + j, bv); indices[(size_t)t * k + j] = bi; }
#pragma unroll
        for (int i = 0; i < MAXPL; ++i) if (lane + 32 * i == bi) loc[i] = -1e30f;
    }
}

// softmax_topk_*_kernel topk_kernels.cuh:192-256, in place, one thread per token
template <int DT>
__global__ void moe_softmax_topk_kernel(void* __restrict__ values, int T, int k) {
    griddep_wait();
    griddep_launch();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    float mx = -INFINITY, sum = 0.f;
    for (int i = 0; i < k; ++i) mx = fmaxf(mx, mld<DT>(values, (size_t)t * k + i));
    for (int i = 0; i < k; ++i) sum += __expf(mld<DT>(values, (size_t)t * k + i) - mx);
    float inv = 1.0f / sum;
    for (int i = 0; i < k; ++i) {
        float e = __expf(mld<DT>(values, (size_t)t * k + i) - mx);
        mst<DT>(values, (size_t)t * k + i, e * inv);
    }
}

// count/offsets/permute/reverse in one single-CTA kernel (T*k <= 512 at decode), ascending order.
__global__ void moe_permutation_kernel(const int32_t* __restrict__ expert_indices, int32_t* __restrict__ counts,
                                       int32_t* __restrict__ offsets, int32_t* __restrict__ permute,
                                       int32_t* __restrict__ reverse, int total, int E) {
    griddep_wait();
    griddep_launch();
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        int c = 0;
        for (int i = 0; i < total; ++i) c += (expert_indices[i] == e);
        counts[e] = c;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int off = 0;
        for (int e = 0; e < E; ++e) { offsets[e] = off; off += counts[e]; }
        offsets[E] = off;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        int p = offsets[e];
        for (int i = 0; i < total; ++i)
            if (expert_indices[i] == e) { permute[p] = i; reverse[i] = p; ++p; }
    }
}

// expand_expert_offsets_kernel moe_kernels.cuh:236-259
__global__ void moe_expand_offsets_kernel(const int32_t* __restrict__ offsets, int32_t* __restrict__ row_ids, int E,
                                          int m_total) {
    griddep_wait();
    griddep_launch();
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m_total) return;
    int lo = 0, hi = E;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (offsets[mid + 1] <= r) lo = mid + 1; else hi = mid; }
    row_ids[r] = lo;
}

// gather_hidden_states_* permute_kernels.cuh:95-182: gathered[p] = hidden[permute[p] / k]
__global__ void moe_gather_kernel(const uint16_t* __restrict__ hidden, const int32_t* __restrict__ permute,
                                  uint16_t* __restrict__ gathered, int rows, int H, int k) {
    griddep_wait();
    griddep_launch();
    for (int p = blockIdx.x; p < rows; p += gridDim.x) {
        const uint16_t* src = hidden + (size_t)(permute[p] / k) * H;
        for (int i = threadIdx.x; i < H; i += blockDim.x) gathered[(size_t)p * H + i] = src[i];
    }
}

// scatter_with_reverse_perm_kernel permute_kernels.cuh:218-242: f32 sum over slots 0..k-1
__global__ void moe_scatter_kernel(const uint16_t* __restrict__ expert_out, const uint16_t* __restrict__ weights,
                                   const int32_t* __restrict__ reverse, uint16_t* __restrict__ out, int T, int H, int k) {
    griddep_wait();
    griddep_launch();
    for (int t = blockIdx.x; t < T; t += gridDim.x)
        for (int h = threadIdx.x; h < H; h += blockDim.x) {
            float sum = 0.f;
            for (int s = 0; s < k; ++s) {
                int f = t * k + s;
                sum = fmaf(bf16_bits_to_f32(weights[f]), bf16_bits_to_f32(expert_out[(size_t)reverse[f] * H + h]), sum);
            }
            out[(size_t)t * H + h] = f32_to_bf16_bits(sum);
        }
}

// Lane-partial dot products of one bf16 activation row with NB FP8 weight rows (block scales per
// 128 k), then a warp sum: 128-bit weight loads, exact f32 products, lane-strided 16-k chunks.  Every
// expert GEMV below goes through this one function, so fused and unfused forms agree bit for bit.
template <int NB, bool CLEAN>
__device__ __forceinline__ void fp8_rows_dot(const uint16_t* __restrict__ arow, const uint8_t* const (&brow)[NB],
                                             const uint16_t* const (&srow)[NB], int K, int lane, float (&acc)[NB]) {
#pragma unroll
    for (int b = 0; b < NB; ++b) acc[b] = 0.f;
    if ((K & 15) == 0) {
#pragma unroll 4
        for (int c = lane; c < (K >> 4); c += 32) {
            uint4 a0 = ld_act16(arow + c * 16);
            uint4 a1 = ld_act16(arow + c * 16 + 8);
            const uint32_t aa[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                uint4 w = ldg_stream(brow[b] + (size_t)c * 16);
                float sc = bf16_bits_to_f32(srow[b][c >> 3]);
                const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
                float part = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint32_t h0 = e4m3x2_to_f16x2_lo(ww[q]), h1 = e4m3x2_to_f16x2_hi(ww[q]);
                    float2 f0 = __half22float2(*reinterpret_cast<__half2*>(&h0));
                    float2 f1 = __half22float2(*reinterpret_cast<__half2*>(&h1));
                    // codes 0x7F / 0xFF are +-480 in the reference's tables (fp8.cuh:89,121), NaN for the hardware
                    // conversion: a byte whose low seven bits are all ones carries into bit 7 when 1 is added
                    // (CLEAN: the tensor was scanned once on the host side and holds neither code -- gemv_fp8.cu::fp8_codes_clean)
                    if (!CLEAN && (((ww[q] & 0x7f7f7f7fu) + 0x01010101u) & 0x80808080u)) {
                        f0.x = fp8_e4m3_to_f32(ww[q] & 0xffu); f0.y = fp8_e4m3_to_f32((ww[q] >> 8) & 0xffu);
                        f1.x = fp8_e4m3_to_f32((ww[q] >> 16) & 0xffu); f1.y = fp8_e4m3_to_f32(ww[q] >> 24);
                    }
                    part = fmaf(bf16lo(aa[2 * q]), f0.x, part);
                    part = fmaf(bf16hi(aa[2 * q]), f0.y, part);
                    part = fmaf(bf16lo(aa[2 * q + 1]), f1.x, part);
                    part = fmaf(bf16hi(aa[2 * q + 1]), f1.y, part);
                }
                acc[b] = fmaf(part, sc, acc[b]);
            }
        }
    } else {
        for (int k = lane; k < K; k += 32)
#pragma unroll
            for (int b = 0; b < NB; ++b)
                acc[b] = fmaf(bf16_bits_to_f32(arow[k]), fp8_e4m3_to_f32(brow[b][k]) * bf16_bits_to_f32(srow[b][k >> 7]), acc[b]);
    }
#pragma unroll
    for (int b = 0; b < NB; ++b) acc[b] = warp_sum(acc[b]);
}

// Expert-grouped FP8 GEMV: row m uses expert row_expert_ids[m] of B[E,N,K], S[E,N/128,K/128].
// One warp per (m, n): 128-bit weight loads, exact f32 products (bf16 activation x decoded
// e4m3), scale per 128-k block.  K % 16 == 0 fast path, scalar tail otherwise.
template <bool CLEAN>
__global__ void moe_grouped_gemv_fp8_kernel(const uint16_t* __restrict__ A, const uint8_t* __restrict__ B,
                                            const uint16_t* __restrict__ S, uint16_t* __restrict__ C,
                                            const int32_t* __restrict__ row_ids, int M, int K, int N,
                                            int a_stride, int expert_offset, int num_local) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31;
    const long gw = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= (long)M * N) return;
    const int n = (int)(gw % N), m = (int)(gw / N);
    const int e = row_ids[m] - expert_offset;
    if (e < 0 || e >= num_local) {             // expert lives on another rank: this rank contributes zero
        if (lane == 0) C[(size_t)m * N + n] = 0;
        return;
    }
    const int SN = (N + 127) >> 7, SK = (K + 127) >> 7;
    const uint8_t* brow[1] = {B + ((size_t)e * N + n) * K};
    const uint16_t* srow[1] = {S + ((size_t)e * SN + (n >> 7)) * SK};
    float acc1[1];
    fp8_rows_dot<1, CLEAN>(A + (size_t)m * a_stride, brow, srow, K, lane, acc1);
    float acc = acc1[0];
    if (lane == 0) C[(size_t)m * N + n] = f32_to_bf16_bits(acc);
}

// One-token routing inside a warp: top-k by k passes of strict '>' argmax (lowest index wins ties) over E <= 512
// router logits, then softmax over the selected values in slot order -- the arithmetic of topk_with_indices +
// softmax_topk (topk_kernels.cuh:111-147,222-256), statement for statement.  Lane j < k returns expert j and its
// bf16 weight bits.
__device__ __forceinline__ void moe_route_warp(const uint16_t* __restrict__ logits, int E, int k, int lane, int& my_idx,
                                               uint16_t& my_w) {
    constexpr int MAXPL = 16;
    float loc[MAXPL];
#pragma unroll
    for (int i = 0; i < MAXPL; ++i) { int e = lane + 32 * i; loc[i] = (e < E) ? bf16_bits_to_f32(logits[e]) : -INFINITY; }
    float mine = -INFINITY;                       // lane j keeps selected value j (k <= 32)
    my_idx = 0;
    for (int j = 0; j < k; ++j) {
        float bv = -1e30f;
        int bi = 0x7fffffff;
#pragma unroll
        for (int i = 0; i < MAXPL; ++i) { int e = lane + 32 * i; if (e < E && loc[i] > bv) { bv = loc[i]; bi = e; } }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (bi == 0x7fffffff) bi = 0;
        if (lane == j) { mine = bf16_bits_to_f32(f32_to_bf16_bits(bv)); my_idx = bi; }
#pragma unroll
        for (int i = 0; i < MAXPL; ++i) if (lane + 32 * i == bi) loc[i] = -1e30f;
    }
    float mx = mine;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float ex = __expf(mine - mx);
    float sum = 0.f;
    for (int i = 0; i < k; ++i) sum += __shfl_sync(0xffffffffu, ex, i);     // slot order, as the serial kernel
    const float inv = 1.0f / sum;
    my_w = f32_to_bf16_bits(ex * inv);
}

// ---------------------------------------------------------------------------------------------
// Expert-grouped NVF4 GEMV (net-new: the reference's MoE path is FP8-only, layers/moe.py:86-151).
// Row r of A[R,K] uses expert e = row_expert_ids[r] - expert_offset of B_data[E,K/2,N] /
// B_scale[E,K/32,N] (the reference NVF4 layout per expert, nvf4_kernels.cu:79-200); rows whose
// expert is not held by this rank produce zeros (their value arrives through the all-reduce).
// CTA = 16 warps over one row x 128 columns: lane owns 4 consecutive columns (32-bit loads, one
// 128-byte line per warp per byte-row), warp w owns scale blocks w, w+16, ... of K; partial sums
// meet in shared memory and are folded in warp order (deterministic).
constexpr int kMoeNvf4Warps = 16;   // 512 threads: 128 registers per thread (32 warps capped the kernel at 64 and spilled ~500 B per thread)
// NMAT = 1: C[r, :] = A[r, :] . B[e]      NMAT = 2 (decode gate/up pair): C[r, :] = bf16(bf16(silu(g)) * u) with
// g = bf16(A . Bg[e]), u = bf16(A . Bu[e]) -- the rounding points of gate GEMV -> silu -> up GEMV -> mul_inplace.
// (A 32-column CTA form with byte loads and a form that recomputed the routing in every warp were measured slower on
// B200 and removed; numbers in profiles/r01_bench_qwen3-30b-a3b_*.json and DESIGN.md.)
template <int NMAT>
__global__ void __launch_bounds__(kMoeNvf4Warps * 32)
moe_grouped_gemv_nvf4_kernel(const uint16_t* __restrict__ A, const uint8_t* __restrict__ Bd0,
                             const uint8_t* __restrict__ Bs0, const uint8_t* __restrict__ Bd1,
                             const uint8_t* __restrict__ Bs1, uint16_t* __restrict__ C,
                             const int32_t* __restrict__ row_expert_ids, int K, int N, int a_stride,
                             int expert_offset, int num_local, float* __restrict__ split_ws,
                             unsigned* __restrict__ split_cnt) {
    extern __shared__ __align__(16) unsigned char moe_smem[];
    __shared__ unsigned s_ticket;
    float* a_s = reinterpret_cast<float*>(moe_smem);                  // [K]
    constexpr int CPL = 4;                                            // columns per lane: CTA = 128 columns, 32-bit loads
    constexpr int TILE = 32 * CPL;
    float* red = a_s + K;                                             // [NMAT][warps][TILE]
    griddep_wait();
    griddep_launch();
    const int r = blockIdx.y, col0 = blockIdx.x * TILE;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int e = row_expert_ids[r] - expert_offset;
    const int c = col0 + lane * CPL;
    if (e < 0 || e >= num_local) {
        if (warp == 0 && c < N) {
#pragma unroll
            for (int j = 0; j < CPL; ++j) if (c + j < N) C[(size_t)r * N + c + j] = 0;
        }
        return;
    }
    for (int k = threadIdx.x; k < K; k += blockDim.x) a_s[k] = bf16_bits_to_f32(A[(size_t)r * a_stride + k]);
    __syncthreads();
    const uint8_t* bdp[2] = {Bd0 + (size_t)e * (K >> 1) * N, NMAT == 2 ? Bd1 + (size_t)e * (K >> 1) * N : nullptr};
    const uint8_t* bsp[2] = {Bs0 + (size_t)e * ((K + 31) >> 5) * N, NMAT == 2 ? Bs1 + (size_t)e * ((K + 31) >> 5) * N : nullptr};
    float acc[NMAT][CPL];
#pragma unroll
    for (int m = 0; m < NMAT; ++m)
#pragma unroll
        for (int j = 0; j < CPL; ++j) acc[m][j] = 0.f;
    const int KB = K >> 5;
    // K split over gridDim.z CTAs (few rows x few column tiles would leave most SMs idle): CTA z owns a contiguous
    // range of scale blocks; the per-CTA column sums meet in a workspace and the last-arriving CTA folds them in
    // z order (deterministic) before the epilogue
    const int KS = (int)gridDim.z, kz = (int)blockIdx.z;
    const int kper = (KB + KS - 1) / KS;
    const int blk0 = kz * kper, blk1 = min(KB, blk0 + kper);
    const bool full = CPL == 4 && (c + CPL - 1 < N) && ((N & (CPL - 1)) == 0);
    for (int blk = blk0 + warp; blk < blk1; blk += kMoeNvf4Warps) {
        if (full) {
            uint32_t w[NMAT][16], sw[NMAT];
#pragma unroll
            for (int m = 0; m < NMAT; ++m) {
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    w[m][i] = CPL == 4 ? __ldg(reinterpret_cast<const uint32_t*>(bdp[m] + (size_t)(blk * 16 + i) * N + c))
                                       : (uint32_t)__ldg(bdp[m] + (size_t)(blk * 16 + i) * N + c);
                sw[m] = CPL == 4 ? __ldg(reinterpret_cast<const uint32_t*>(bsp[m] + (size_t)blk * N + c))
                                 : (uint32_t)__ldg(bsp[m] + (size_t)blk * N + c);
            }
#pragma unroll
            for (int m = 0; m < NMAT; ++m) {
                float part[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float2 a = *reinterpret_cast<const float2*>(a_s + blk * 32 + 2 * i);
                    uint32_t h;
                    float2 f;
                    h = e2m1x2_to_f16x2<0>(w[m][i]); f = __half22float2(*reinterpret_cast<__half2*>(&h));
                    part[0] = fmaf(a.x, f.x, part[0]); part[0] = fmaf(a.y, f.y, part[0]);
                    if (CPL == 4) {
                        h = e2m1x2_to_f16x2<1>(w[m][i]); f = __half22float2(*reinterpret_cast<__half2*>(&h));
                        part[1] = fmaf(a.x, f.x, part[1]); part[1] = fmaf(a.y, f.y, part[1]);
                        h = e2m1x2_to_f16x2<2>(w[m][i]); f = __half22float2(*reinterpret_cast<__half2*>(&h));
                        part[2] = fmaf(a.x, f.x, part[2]); part[2] = fmaf(a.y, f.y, part[2]);
                        h = e2m1x2_to_f16x2<3>(w[m][i]); f = __half22float2(*reinterpret_cast<__half2*>(&h));
                        part[3] = fmaf(a.x, f.x, part[3]); part[3] = fmaf(a.y, f.y, part[3]);
                    }
                }
#pragma unroll
                for (int j = 0; j < CPL; ++j) acc[m][j] = fmaf(part[j], ue4m3_to_f32((sw[m] >> (8 * j)) & 0xff), acc[m][j]);
            }
        } else {
#pragma unroll
            for (int m = 0; m < NMAT; ++m) {
                float part[4] = {0.f, 0.f, 0.f, 0.f};
                for (int j = 0; j < CPL; ++j) {
                    if (c + j >= N) break;
                    uint32_t b[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) b[i] = __ldg(bdp[m] + (size_t)(blk * 16 + i) * N + c + j);
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        part[j] = fmaf(a_s[blk * 32 + 2 * i], e2m1_to_f32(b[i] & 0xf), part[j]);
                        part[j] = fmaf(a_s[blk * 32 + 2 * i + 1], e2m1_to_f32(b[i] >> 4), part[j]);
                    }
                    acc[m][j] = fmaf(part[j], ue4m3_to_f32(bsp[m][(size_t)blk * N + c + j]), acc[m][j]);
                }
            }
        }
    }
#pragma unroll
    for (int m = 0; m < NMAT; ++m)
#pragma unroll
        for (int j = 0; j < CPL; ++j) red[(m * kMoeNvf4Warps + warp) * TILE + lane * CPL + j] = acc[m][j];
    __syncthreads();
    float s[NMAT];
    if (threadIdx.x < TILE) {
#pragma unroll
        for (int m = 0; m < NMAT; ++m) {
            s[m] = 0.f;
#pragma unroll 8
            for (int w = 0; w < kMoeNvf4Warps; ++w) s[m] += red[(m * kMoeNvf4Warps + w) * TILE + threadIdx.x];
        }
    }
    if (KS > 1) {
        const size_t cell = (size_t)r * gridDim.x + blockIdx.x;
        float* wsp = split_ws + cell * KS * NMAT * TILE;
        if (threadIdx.x < TILE) {
#pragma unroll
            for (int m = 0; m < NMAT; ++m) __stcg(wsp + ((size_t)kz * NMAT + m) * TILE + threadIdx.x, s[m]);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            s_ticket = atomicAdd(split_cnt + cell, 1u);
            __threadfence();
        }
        __syncthreads();
        if (s_ticket != (unsigned)(KS - 1)) return;
        if (threadIdx.x < TILE) {
#pragma unroll
            for (int m = 0; m < NMAT; ++m) {
                float t = 0.f;
                for (int zz = 0; zz < KS; ++zz) t += __ldcg(wsp + ((size_t)zz * NMAT + m) * TILE + threadIdx.x);
                s[m] = t;
            }
        }
        if (threadIdx.x == 0) split_cnt[cell] = 0;
    }
    if (threadIdx.x < TILE && col0 + threadIdx.x < N) {
        uint16_t o;
        if (NMAT == 2) {
            const float g = bf16_bits_to_f32(f32_to_bf16_bits(s[0]));
            const float u = bf16_bits_to_f32(f32_to_bf16_bits(s[NMAT - 1]));
            o = f32_to_bf16_bits(bf16_bits_to_f32(f32_to_bf16_bits(silu_ref(g))) * u);
        } else {
            o = f32_to_bf16_bits(s[0]);
        }
        C[(size_t)r * N + col0 + threadIdx.x] = o;
    }
}

// ---------------------------------------------------------------------------------------------
// Decode (one token) forms.  With T == 1 the permutation is the identity up to the order of the k
// rows, and the combine sums in SLOT order whatever the row order is, so row r can simply be slot r:
// routing collapses to one warp and gather / permutation / expand disappear.  Arithmetic is that of
// the generic kernels above, statement for statement (tests compare the two bit for bit).

// top-k (strict '>', lowest index wins) + softmax over the selected logits + row ids, one warp.
__global__ void moe_route_t1_kernel(const uint16_t* __restrict__ logits, uint16_t* __restrict__ weights,
                                    int32_t* __restrict__ indices, int32_t* __restrict__ row_ids, int E, int k) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x;
    int idx;
    uint16_t w;
    moe_route_warp(logits, E, k, lane, idx, w);
    if (lane < k) { indices[lane] = idx; row_ids[lane] = idx; weights[lane] = w; }
}

// gate and up GEMVs of the k routed experts + silu + multiply in one launch: one warp per (slot, n)
// reads x once for both weight rows.  act = bf16(bf16(silu(bf16 gate)) * bf16 up), the three-launch form.
template <bool CLEAN>
__global__ void moe_gateup_fp8_t1_kernel(const uint16_t* __restrict__ x, const uint8_t* __restrict__ Bg,
                                         const uint16_t* __restrict__ Sg, const uint8_t* __restrict__ Bu,
                                         const uint16_t* __restrict__ Su, uint16_t* __restrict__ act,
                                         const int32_t* __restrict__ row_ids, int R, int K, int N, int expert_offset,
                                         int num_local) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31;
    const long gw = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= (long)R * N) return;
    const int n = (int)(gw % N), m = (int)(gw / N);
    const int e = row_ids[m] - expert_offset;
    if (e < 0 || e >= num_local) {
        if (lane == 0) act[(size_t)m * N + n] = 0;
        return;
    }
    const int SN = (N + 127) >> 7, SK = (K + 127) >> 7;
    const size_t wo = ((size_t)e * N + n) * K, so = ((size_t)e * SN + (n >> 7)) * SK;
    const uint8_t* brow[2] = {Bg + wo, Bu + wo};
    const uint16_t* srow[2] = {Sg + so, Su + so};
    float acc[2];
    fp8_rows_dot<2, CLEAN>(x, brow, srow, K, lane, acc);
    if (lane == 0) {
        const float g = bf16_bits_to_f32(f32_to_bf16_bits(acc[0]));
        const float u = bf16_bits_to_f32(f32_to_bf16_bits(acc[1]));
        const float sg = bf16_bits_to_f32(f32_to_bf16_bits(silu_ref(g)));
        act[(size_t)m * N + n] = f32_to_bf16_bits(sg * u);
    }
}

// down GEMV of the k routed experts + weighted combine in slot order (+ residual): CTA = cols x k
// warps, warp (c, s) owns expert row n0+c of slot s; the k rounded results meet in shared memory.
template <bool CLEAN>
__global__ void moe_down_combine_fp8_t1_kernel(const uint16_t* __restrict__ act, const uint8_t* __restrict__ Bd,
                                               const uint16_t* __restrict__ Sd, const uint16_t* __restrict__ weights,
                                               const int32_t* __restrict__ row_ids, const uint16_t* __restrict__ residual,
                                               uint16_t* __restrict__ out, int k, int K, int N, int cols,
                                               int expert_offset, int num_local) {
    __shared__ float part_s[32];
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = warp / k, sl = warp - c * k;
    const int n = blockIdx.x * cols + c;
    float d = 0.f;
    if (n < N) {
        const int e = row_ids[sl] - expert_offset;
        if (e >= 0 && e < num_local) {
            const int SN = (N + 127) >> 7, SK = (K + 127) >> 7;
            const uint8_t* brow[1] = {Bd + ((size_t)e * N + n) * K};
            const uint16_t* srow[1] = {Sd + ((size_t)e * SN + (n >> 7)) * SK};
            float acc[1];
            fp8_rows_dot<1, CLEAN>(act + (size_t)sl * K, brow, srow, K, lane, acc);
            d = bf16_bits_to_f32(f32_to_bf16_bits(acc[0]));
        }
    }
    if (lane == 0) part_s[warp] = d;
    __syncthreads();
    if (threadIdx.x < cols) {
        const int nn = blockIdx.x * cols + threadIdx.x;
        if (nn < N) {
            float sum = 0.f;
            for (int s2 = 0; s2 < k; ++s2) sum = fmaf(bf16_bits_to_f32(weights[s2]), part_s[threadIdx.x * k + s2], sum);
            uint16_t o = f32_to_bf16_bits(sum);
            if (residual) o = f32_to_bf16_bits(bf16_bits_to_f32(o) + bf16_bits_to_f32(residual[nn]));
            out[nn] = o;
        }
    }
}

// scatter with identity row order (+ residual) for the formats without a fused down kernel
__global__ void moe_combine_t1_kernel(const uint16_t* __restrict__ expert_out, const uint16_t* __restrict__ weights,
                                      const uint16_t* __restrict__ residual, uint16_t* __restrict__ out, int H, int k) {
    griddep_wait();
    griddep_launch();
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    float sum = 0.f;
    for (int s = 0; s < k; ++s)
        sum = fmaf(bf16_bits_to_f32(weights[s]), bf16_bits_to_f32(expert_out[(size_t)s * H + h]), sum);
    uint16_t o = f32_to_bf16_bits(sum);
    if (residual) o = f32_to_bf16_bits(bf16_bits_to_f32(o) + bf16_bits_to_f32(residual[h]));
    out[h] = o;
}

static inline unsigned nblk(size_t n, int bs) { return (unsigned)((n + bs - 1) / bs); }

static float* g_moe_split_ws = nullptr;
static unsigned* g_moe_split_cnt = nullptr;

template <int NMAT>
static int launch_moe_nvf4(const void* A, int a_stride, const void* Bd0, const void* Bs0, const void* Bd1,
                           const void* Bs1, void* C, const int32_t* row_expert_ids, int M, int K, int N,
                           int expert_offset, int num_local_experts, void* stream) {
    constexpr int TILE = 128;
    size_t smem = (size_t)K * 4 + (size_t)NMAT * kMoeNvf4Warps * TILE * 4;
    PGK_REQUIRE(smem <= 200 * 1024, "moe_grouped_gemv_nvf4 failed: K=%d too large", K);
    static bool attr = false;
    if (!attr) {
        PGK_CUDA(cudaFuncSetAttribute(moe_grouped_gemv_nvf4_kernel<NMAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr = true;
    }
    // K split: when rows x column tiles cover less than half of the SMs, spread the scale blocks of every cell over up
    // to 4 CTAs (each keeps >= 16 blocks = one per warp).  The workspace is allocated on the first call made outside
    // graph capture; until then (and when it would not fit) the kernel runs unsplit.  PGK_MOE_NVF4_KS=1 disables.
    float*& ws = g_moe_split_ws;          // shared by all instantiations: allocated by whichever runs first (prefill, eager)
    unsigned*& cnt = g_moe_split_cnt;
    static const int ks_max = [] { const char* e = getenv("PGK_MOE_NVF4_KS"); int v = e ? atoi(e) : 4; return v < 1 ? 1 : (v > 8 ? 8 : v); }();
    constexpr size_t kCells = 128, kWsFloats = kCells * 8 * 2 * 128;       // split only when cells * 2 <= SM count
    const int tiles = (N + TILE - 1) / TILE;
    if (!ws) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing((cudaStream_t)stream, &cs);
        if (cs == cudaStreamCaptureStatusNone) {
            PGK_CUDA(cudaMalloc(&ws, kWsFloats * sizeof(float)));
            PGK_CUDA(cudaMalloc(&cnt, kCells * sizeof(unsigned)));
            PGK_CUDA(cudaMemset(cnt, 0, kCells * sizeof(unsigned)));
        }
    }
    int ks = 1;
    const long cells = (long)tiles * M;
    if (ws && cells * 2 <= sm_count() && cells <= (long)kCells) {
        ks = (int)(sm_count() / cells);
        if (ks > ks_max) ks = ks_max;
        const int kb = K >> 5;
        while (ks > 1 && kb / ks < kMoeNvf4Warps) --ks;
    }
    PGK_CUDA(launch(moe_grouped_gemv_nvf4_kernel<NMAT>, dim3(tiles, M, ks), dim3(kMoeNvf4Warps * 32), smem,
                    (cudaStream_t)stream, (const uint16_t*)A, (const uint8_t*)Bd0, (const uint8_t*)Bs0, (const uint8_t*)Bd1,
                    (const uint8_t*)Bs1, (uint16_t*)C, row_expert_ids, K, N, a_stride, expert_offset, num_local_experts,
                    ws, cnt));
    return 0;
}
}  // namespace pgk

using namespace pgk;
#define ST(s) ((cudaStream_t)(s))

// One scan per stacked expert tensor (cached by pointer): tensors without the codes 0x7F / 0xFF take the kernels
// without the per-word test for them.  An unknown expert count (0x7fffffff) keeps the exact kernels.
static int moe_fp8_clean(const void* B_enk, int num_local, int N, int K, void* stream, bool* clean) {
    *clean = false;
    if (num_local <= 0 || num_local == 0x7fffffff) return 0;
    return pgk::fp8_codes_clean(B_enk, (size_t)num_local * N * K, ST(stream), clean);
}

extern "C" {

int pgk_moe_topk(const void* logits, void* values, int32_t* indices, int num_tokens, int num_experts, int k,
                 int dtype, void* stream) {
    PGK_REQUIRE(num_experts >= 1 && num_experts <= 512 && k >= 1 && k <= num_experts,
                "moe_topk_with_indices failed: unsupported num_experts=%d k=%d", num_experts, k);
    if (num_tokens <= 0) return 0;
    unsigned g = nblk((size_t)num_tokens * 32, 128);
    if (dtype == 0) PGK_CUDA(launch(moe_topk_kernel<0>, dim3(g), dim3(128), 0, ST(stream), logits, values, indices, num_tokens, num_experts, k));
    else if (dtype == 1) PGK_CUDA(launch(moe_topk_kernel<1>, dim3(g), dim3(128), 0, ST(stream), logits, values, indices, num_tokens, num_experts, k));
    else if (dtype == 2) PGK_CUDA(launch(moe_topk_kernel<2>, dim3(g), dim3(128), 0, ST(stream), logits, values, indices, num_tokens, num_experts, k));
    else { set_error("moe_topk_with_indices failed: unsupported dtype code %d", dtype); return 1; }
    return 0;
}
int pgk_moe_softmax_topk(void* values, int num_tokens, int k, int dtype, void* stream) {
    if (num_tokens <= 0) return 0;
    unsigned g = nblk(num_tokens, 128);
    if (dtype == 0) PGK_CUDA(launch(moe_softmax_topk_kernel<0>, dim3(g), dim3(128), 0, ST(stream), values, num_tokens, k));
    else if (dtype == 1) PGK_CUDA(launch(moe_softmax_topk_kernel<1>, dim3(g), dim3(128), 0, ST(stream), values, num_tokens, k));
    else if (dtype == 2) PGK_CUDA(launch(moe_softmax_topk_kernel<2>, dim3(g), dim3(128), 0, ST(stream), values, num_tokens, k));
    else { set_error("moe_softmax_topk failed: unsupported dtype code %d", dtype); return 1; }
    return 0;
}
int pgk_moe_compute_permutation(const int32_t* expert_indices, int32_t* expert_counts, int32_t* expert_offsets,
                                int32_t* permute_indices, int32_t* reverse_perm, int num_tokens, int num_experts,
                                int k, void* stream) {
    PGK_CUDA(launch(moe_permutation_kernel, dim3(1), dim3(256), 0, ST(stream), expert_indices, expert_counts,
                    expert_offsets, permute_indices, reverse_perm, num_tokens * k, num_experts));
    return 0;
}
int pgk_moe_expand_expert_offsets(const int32_t* expert_offsets, int32_t* row_expert_ids, int num_experts,
                                  int m_total, void* stream) {
    if (m_total <= 0) return 0;
    PGK_CUDA(launch(moe_expand_offsets_kernel, dim3(nblk(m_total, 128)), dim3(128), 0, ST(stream), expert_offsets,
                    row_expert_ids, num_experts, m_total));
    return 0;
}
int pgk_moe_gather(const void* hidden, const int32_t* permute_indices, void* gathered, int num_tokens,
                   int hidden_size, int k, void* stream) {
    int rows = num_tokens * k;
    if (rows <= 0) return 0;
    PGK_CUDA(launch(moe_gather_kernel, dim3(rows), dim3(256), 0, ST(stream), (const uint16_t*)hidden, permute_indices,
                    (uint16_t*)gathered, rows, hidden_size, k));
    return 0;
}
int pgk_moe_scatter(const void* expert_out, const void* router_weights, const int32_t* reverse_perm, void* output,
                    int num_tokens, int hidden_size, int k, void* stream) {
    if (num_tokens <= 0) return 0;
    PGK_CUDA(launch(moe_scatter_kernel, dim3(num_tokens), dim3(256), 0, ST(stream), (const uint16_t*)expert_out,
                    (const uint16_t*)router_weights, reverse_perm, (uint16_t*)output, num_tokens, hidden_size, k));
    return 0;
}
int pgk_moe_grouped_gemv_fp8_ep(const void* A, const void* B_enk, const void* S, void* C, const int32_t* row_expert_ids,
                                int M, int K, int N, int expert_offset, int num_local_experts, void* stream) {
    return pgk_moe_grouped_gemv_fp8_ex(A, K, B_enk, S, C, row_expert_ids, M, K, N, expert_offset, num_local_experts, stream);
}
int pgk_moe_grouped_gemv_fp8_ex(const void* A, int a_stride, const void* B_enk, const void* S, void* C,
                                const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                                int num_local_experts, void* stream) {
    if (M <= 0) return 0;
    PGK_REQUIRE(a_stride == 0 || a_stride >= K, "moe_grouped_gemv_fp8 failed: a_stride=%d must be 0 (broadcast) or >= K", a_stride);
    size_t warps = (size_t)M * N;
    bool clean;
    if (int rc = moe_fp8_clean(B_enk, num_local_experts, N, K, stream, &clean)) return rc;
    auto kern = clean ? moe_grouped_gemv_fp8_kernel<true> : moe_grouped_gemv_fp8_kernel<false>;
    PGK_CUDA(launch(kern, dim3(nblk(warps * 32, 256)), dim3(256), 0, ST(stream),
                    (const uint16_t*)A, (const uint8_t*)B_enk, (const uint16_t*)S, (uint16_t*)C, row_expert_ids, M, K, N,
                    a_stride, expert_offset, num_local_experts));
    return 0;
}
int pgk_moe_grouped_gemv_fp8(const void* A, const void* B_enk, const void* S, void* C, const int32_t* row_expert_ids,
                             int M, int K, int N, void* stream) {
    return pgk_moe_grouped_gemv_fp8_ep(A, B_enk, S, C, row_expert_ids, M, K, N, 0, 0x7fffffff, stream);
}

int pgk_moe_grouped_gemv_nvf4(const void* A, const void* B_data, const void* B_scale, void* C,
                              const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                              int num_local_experts, void* stream) {
    return pgk_moe_grouped_gemv_nvf4_ex(A, K, B_data, B_scale, C, row_expert_ids, M, K, N, expert_offset,
                                        num_local_experts, stream);
}
int pgk_moe_grouped_gemv_nvf4_ex(const void* A, int a_stride, const void* B_data, const void* B_scale, void* C,
                                 const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                                 int num_local_experts, void* stream) {
    if (M <= 0) return 0;
    PGK_REQUIRE(a_stride == 0 || a_stride >= K, "moe_grouped_gemv_nvf4 failed: a_stride=%d must be 0 (broadcast) or >= K", a_stride);
    PGK_REQUIRE(K % 32 == 0 && K >= 32, "moe_grouped_gemv_nvf4 failed: K=%d must be a multiple of 32", K);
    return launch_moe_nvf4<1>(A, a_stride, B_data, B_scale, nullptr, nullptr, C, row_expert_ids, M, K, N, expert_offset,
                              num_local_experts, stream);
}
int pgk_moe_gateup_nvf4_t1(const void* x, const void* Bg_data, const void* Bg_scale, const void* Bu_data,
                           const void* Bu_scale, void* act, const int32_t* row_expert_ids, int k, int K, int N,
                           int expert_offset, int num_local_experts, void* stream) {
    if (k <= 0) return 0;
    PGK_REQUIRE(K % 32 == 0 && K >= 32, "moe_gateup_nvf4_t1 failed: K=%d must be a multiple of 32", K);
    return launch_moe_nvf4<2>(x, 0, Bg_data, Bg_scale, Bu_data, Bu_scale, act, row_expert_ids, k, K, N, expert_offset,
                              num_local_experts, stream);
    return 0;
}

/* ---- decode (one token) forms: same results as the generic pipeline, fewer launches ---- */
int pgk_moe_route_t1(const void* logits, void* weights, int32_t* indices, int32_t* row_expert_ids, int num_experts,
                     int k, void* stream) {
    PGK_REQUIRE(num_experts >= 1 && num_experts <= 512 && k >= 1 && k <= num_experts && k <= 32,
                "moe_route_t1 failed: unsupported num_experts=%d k=%d", num_experts, k);
    PGK_CUDA(launch(moe_route_t1_kernel, dim3(1), dim3(32), 0, ST(stream), (const uint16_t*)logits, (uint16_t*)weights,
                    indices, row_expert_ids, num_experts, k));
    return 0;
}
int pgk_moe_gateup_fp8_t1(const void* x, const void* Bg_enk, const void* Sg, const void* Bu_enk, const void* Su,
                          void* act, const int32_t* row_expert_ids, int k, int K, int N, int expert_offset,
                          int num_local_experts, void* stream) {
    if (k <= 0) return 0;
    size_t warps = (size_t)k * N;
    bool cg, cu;
    if (int rc = moe_fp8_clean(Bg_enk, num_local_experts, N, K, stream, &cg)) return rc;
    if (int rc = moe_fp8_clean(Bu_enk, num_local_experts, N, K, stream, &cu)) return rc;
    auto kern = (cg && cu) ? moe_gateup_fp8_t1_kernel<true> : moe_gateup_fp8_t1_kernel<false>;
    PGK_CUDA(launch(kern, dim3(nblk(warps * 32, 256)), dim3(256), 0, ST(stream), (const uint16_t*)x,
                    (const uint8_t*)Bg_enk, (const uint16_t*)Sg, (const uint8_t*)Bu_enk, (const uint16_t*)Su,
                    (uint16_t*)act, row_expert_ids, k, K, N, expert_offset, num_local_experts));
    return 0;
}
int pgk_moe_down_combine_fp8_t1(const void* act, const void* Bd_enk, const void* Sd, const void* router_weights,
                                const int32_t* row_expert_ids, const void* residual, void* out, int k, int K, int N,
                                int expert_offset, int num_local_experts, void* stream) {
    PGK_REQUIRE(k >= 1 && k <= 32, "moe_down_combine_fp8_t1 failed: unsupported k=%d", k);
    const int cols = k > 16 ? 1 : (16 / k > 4 ? 4 : 16 / k);   // <= 512-thread CTAs (k <= 16), several resident per SM
    bool clean;
    if (int rc = moe_fp8_clean(Bd_enk, num_local_experts, N, K, stream, &clean)) return rc;
    auto kern = clean ? moe_down_combine_fp8_t1_kernel<true> : moe_down_combine_fp8_t1_kernel<false>;
    PGK_CUDA(launch(kern, dim3((N + cols - 1) / cols), dim3(cols * k * 32), 0, ST(stream),
                    (const uint16_t*)act, (const uint8_t*)Bd_enk, (const uint16_t*)Sd, (const uint16_t*)router_weights,
                    row_expert_ids, (const uint16_t*)residual, (uint16_t*)out, k, K, N, cols, expert_offset,
                    num_local_experts));
    return 0;
}
int pgk_moe_combine_t1(const void* expert_out, const void* router_weights, const void* residual, void* out,
                       int hidden_size, int k, void* stream) {
    PGK_CUDA(launch(moe_combine_t1_kernel, dim3(nblk(hidden_size, 256)), dim3(256), 0, ST(stream),
                    (const uint16_t*)expert_out, (const uint16_t*)router_weights, (const uint16_t*)residual,
                    (uint16_t*)out, hidden_size, k));
    return 0;
}

}  // extern "C"
