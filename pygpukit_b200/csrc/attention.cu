// attention.cu -- split-KV flash-decoding over a fixed KV cache (q_len small), GQA-aware.
//
//   out[h, t, :] = softmax_j( Q[h,t,:] . K[kv(h), j, :] * scale ) V[kv(h), j, :],
//   j < context_len - q_len + t + 1
//
// replaces sdpa_causal_fixed_cache(_ptr) -> sdpa_causal_dispatch (sdpa_causal.inl:452-769):
// the naive sdpa_causal_*_kernel (attention_kernels.cuh:32-705, one thread per KV position,
// scores in shared memory, fails past 12k context) and the FP16-only two-kernel
// flash_decoding (flash_decoding.cuh:75-377).
//
// Design (HBM-bound; DESIGN.md "sdpa_decode"):
//   * the cache is read in its UN-expanded [n_kv, max_seq, hd] form: one CTA serves all
//     G = Hq/n_kv query heads of a KV head (in row tiles of RT), so K/V bytes are fetched
//     once per group instead of G times (the reference stores and reads the expanded copy).
//   * split-KV: `splits` CTAs per (kv head, row tile) each take a contiguous chunk computed
//     from the DEVICE-side context length (graph-replay safe), keep an online-softmax state
//     per query row in registers, and write (o, m, l) partials to an f32 workspace; the
//     last-arriving CTA folds them in split order (deterministic) and writes the output.
//   * K/V rows are loaded with 128-bit coalesced loads (16 lanes cover one 256-byte bf16
//     row, two positions per warp iteration, two iterations in flight).
#include "common.cuh"
#include "../../include/pgk_b200.h"
#include <cuda.h>
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>

namespace pgk {

template <int DT> struct KvVec;   // 16-byte vector of EPL elements -> f32
template <> struct KvVec<0> {
    static constexpr int EPL = 4;
    static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
        f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    }
};
template <> struct KvVec<1> {
    static constexpr int EPL = 8;
    static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
    }
};
template <> struct KvVec<2> {
    static constexpr int EPL = 8;
    static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
        f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
        f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
    }
};
template <int DT> __device__ __forceinline__ float ld_elem(const void* p, size_t i);
template <> __device__ __forceinline__ float ld_elem<0>(const void* p, size_t i) { return ((const float*)p)[i]; }
template <> __device__ __forceinline__ float ld_elem<1>(const void* p, size_t i) { return __half2float(((const __half*)p)[i]); }
template <> __device__ __forceinline__ float ld_elem<2>(const void* p, size_t i) { return __bfloat162float(((const __nv_bfloat16*)p)[i]); }
template <int DT> __device__ __forceinline__ void st_elem(void* p, size_t i, float v);
template <> __device__ __forceinline__ void st_elem<0>(void* p, size_t i, float v) { ((float*)p)[i] = v; }
template <> __device__ __forceinline__ void st_elem<1>(void* p, size_t i, float v) { ((__half*)p)[i] = __float2half_rn(v); }
template <> __device__ __forceinline__ void st_elem<2>(void* p, size_t i, float v) { ((__nv_bfloat16*)p)[i] = __float2bfloat16_rn(v); }

__device__ __forceinline__ uint4 ldg_kv(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

struct SdpaParams {
    const void* Q; const void* K; const void* V; void* out;
    int hq, n_kv, q_len, max_seq, context_len;
    const int32_t* ctx_buf;
    float scale_log2;         // scale * log2(e)
    long q_sh, q_st, o_sh, o_st;
    int splits;
    float* ws;                // [(kvh*row_tiles + rt)*splits + s][RT][HD+2]
    unsigned* counters;       // [n_kv*row_tiles]
};

constexpr int kAttnWarps = 8;
constexpr int kAttnMaxSplits = 128;
constexpr int kAttnU = 4;
constexpr int kAttnStages = 2; // K/V tiles in flight per CTA
// KV positions per staged tile: 32 KB (K + V) per stage for 16-bit caches of head_dim 128
template <int DT, int HD> struct AttnTile { static constexpr int TP = (DT == 0 ? 16 : 32) * 256 / HD >= 16 ? (DT == 0 ? 16 : 32) * 256 / HD : 16; };     // KV positions per lane group per batch
constexpr float kNegBig = -1e30f;

template <int DT, int HD, int RT>
__global__ void __launch_bounds__(kAttnWarps * 32, 2)
sdpa_decode_kernel(SdpaParams p) {
    constexpr int EPL = KvVec<DT>::EPL;
    constexpr int LPR = HD / EPL;          // lanes per KV row
    constexpr int PPI = 32 / LPR;          // positions per warp iteration
    constexpr int REC = HD + 2;            // partial record: o[HD], m, l
    static_assert(LPR <= 32 && PPI >= 1, "head_dim too large for one warp row");
    __shared__ float sm_part[kAttnWarps][RT][REC];
    __shared__ unsigned s_ticket;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int li = lane % LPR, slot = lane / LPR;
    const int G = p.hq / p.n_kv;
    const int R = G * p.q_len;
    const int row_tiles = (R + RT - 1) / RT;
    const int split = blockIdx.x, rt = blockIdx.y, kvh = blockIdx.z;

    griddep_wait();
    griddep_launch();
    const int ctx = min(p.ctx_buf ? *p.ctx_buf : p.context_len, p.max_seq);
    int chunk = (ctx + p.splits - 1) / p.splits;
    chunk = (chunk + 31) & ~31;
    const int c0 = split * chunk, c1 = min(ctx, c0 + chunk);

    // query rows of this tile: this lane's EPL-slice, pre-scaled by scale*log2(e)
    float q[RT][EPL];
    int limit[RT];
#pragma unroll
    for (int r = 0; r < RT; ++r) {
        int row = rt * RT + r;
        if (row < R) {
            int g = row / p.q_len, t = row % p.q_len;
            const size_t qo = (size_t)(kvh * G + g) * p.q_sh + (size_t)t * p.q_st + (size_t)li * EPL;
#pragma unroll
            for (int e = 0; e < EPL; ++e) q[r][e] = ld_elem<DT>(p.Q, qo + e) * p.scale_log2;
            limit[r] = ctx - p.q_len + t + 1;
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) q[r][e] = 0.f;
            limit[r] = 0;
        }
    }
    float m[RT], l[RT], o[RT][EPL];
#pragma unroll
    for (int r = 0; r < RT; ++r) {
        m[r] = kNegBig; l[r] = 0.f;
#pragma unroll
        for (int e = 0; e < EPL; ++e) o[r][e] = 0.f;
    }

    constexpr size_t esz = (DT == 0) ? 4 : 2;
    constexpr int TP = AttnTile<DT, HD>::TP;               // KV positions per staged tile
    constexpr uint32_t ROWB = HD * esz;                      // bytes of one K (or V) row
    constexpr uint32_t TILEB = TP * ROWB;                    // bytes of one K (or V) tile
    // K/V tiles are contiguous in the [n_kv, max_seq, hd] cache, so one cp.async.bulk (TMA, UBLKCP) per
    // tensor stages a whole tile in shared memory; kAttnStages tiles are in flight per CTA (>= 64 KB per
    // SM with two resident CTAs), decoupling the HBM latency from the softmax arithmetic.
    extern __shared__ __align__(128) unsigned char kv_smem[];
    __shared__ __align__(8) uint64_t full_bar[kAttnStages];
    const uint8_t* Kg = (const uint8_t*)p.K + (size_t)kvh * p.max_seq * ROWB;
    const uint8_t* Vg = (const uint8_t*)p.V + (size_t)kvh * p.max_seq * ROWB;
    const int ntiles = (c1 > c0) ? (c1 - c0 + TP - 1) / TP : 0;
    auto issue_tile = [&](int t) {                            // one elected thread
        const int st = t % kAttnStages;
        const int t0 = c0 + t * TP;
        const uint32_t bytes = (uint32_t)(min(TP, c1 - t0)) * ROWB;
        unsigned char* ks = kv_smem + (size_t)st * 2 * TILEB;
        mbar_arrive_expect_tx(&full_bar[st], 2 * bytes);
        bulk_copy_g2s(ks, Kg + (size_t)t0 * ROWB, bytes, &full_bar[st]);
        bulk_copy_g2s(ks + TILEB, Vg + (size_t)t0 * ROWB, bytes, &full_bar[st]);
    };
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < kAttnStages; ++i) mbar_init(&full_bar[i], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int t = 0; t < min(ntiles, kAttnStages); ++t) issue_tile(t);

    constexpr int U = kAttnU;
    constexpr int STEP = kAttnWarps * PPI;          // positions per CTA per u-slot
    for (int t = 0; t < ntiles; ++t) {
        const int st = t % kAttnStages;
        const int t0 = c0 + t * TP, t1 = min(c1, t0 + TP);
        mbar_wait(&full_bar[st], (uint32_t)((t / kAttnStages) & 1));
        const unsigned char* ks = kv_smem + (size_t)st * 2 * TILEB + (size_t)li * 16;
        const unsigned char* vs = ks + TILEB;
        // Each lane group (LPR lanes = one KV row) takes U positions per batch; one online-softmax rescale
        // per batch.  The loop bounds are warp-uniform, so the full-mask shuffles are always convergent.
        for (int jw = t0 + warp * PPI; jw < t1; jw += U * STEP) {
            uint4 kq[U], vq[U];
            bool valid[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int j = jw + slot + u * STEP;
                valid[u] = j < t1;
                if (valid[u]) {
                    kq[u] = *reinterpret_cast<const uint4*>(ks + (size_t)(j - t0) * ROWB);
                    vq[u] = *reinterpret_cast<const uint4*>(vs + (size_t)(j - t0) * ROWB);
                } else {
                    kq[u] = make_uint4(0, 0, 0, 0); vq[u] = make_uint4(0, 0, 0, 0);
                }
            }
            float sc[U][RT];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float kf[EPL];
                KvVec<DT>::unpack(kq[u], kf);
                const int j = jw + slot + u * STEP;
#pragma unroll
                for (int r = 0; r < RT; ++r) {
                    float tt = 0.f;
#pragma unroll
                    for (int e = 0; e < EPL; ++e) tt = fmaf(q[r][e], kf[e], tt);
#pragma unroll
                    for (int off = LPR / 2; off > 0; off >>= 1) tt += __shfl_xor_sync(0xffffffffu, tt, off);
                    sc[u][r] = (valid[u] && j < limit[r]) ? tt : kNegBig;
                }
            }
#pragma unroll
            for (int r = 0; r < RT; ++r) {
                float mn = m[r];
#pragma unroll
                for (int u = 0; u < U; ++u) mn = fmaxf(mn, sc[u][r]);
                const float corr = exp2f(m[r] - mn);
                m[r] = mn;
                l[r] *= corr;
#pragma unroll
                for (int e = 0; e < EPL; ++e) o[r][e] *= corr;
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const float pj = (sc[u][r] > 0.5f * kNegBig) ? exp2f(sc[u][r] - mn) : 0.f;
                    sc[u][r] = pj;
                    l[r] += pj;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float vf[EPL];
                KvVec<DT>::unpack(vq[u], vf);
#pragma unroll
                for (int r = 0; r < RT; ++r)
#pragma unroll
                    for (int e = 0; e < EPL; ++e) o[r][e] = fmaf(sc[u][r], vf[e], o[r][e]);
            }
        }
        if (t + kAttnStages < ntiles) {               // refill this stage (CTA-uniform condition)
            __syncthreads();                          // every warp is done reading it
            if (threadIdx.x == 0) { fence_proxy_async(); issue_tile(t + kAttnStages); }
        }
    }
    // merge the PPI slots of the warp (lanes li, li+LPR, ...)
#pragma unroll
    for (int r = 0; r < RT; ++r) {
#pragma unroll
        for (int off = LPR; off < 32; off <<= 1) {
            float m2 = __shfl_xor_sync(0xffffffffu, m[r], off);
            float l2 = __shfl_xor_sync(0xffffffffu, l[r], off);
            float mn = fmaxf(m[r], m2);
            float c1s = exp2f(m[r] - mn), c2s = exp2f(m2 - mn);
            l[r] = l[r] * c1s + l2 * c2s;
#pragma unroll
            for (int e = 0; e < EPL; ++e) {
                float o2 = __shfl_xor_sync(0xffffffffu, o[r][e], off);
                o[r][e] = o[r][e] * c1s + o2 * c2s;
            }
            m[r] = mn;
        }
        if (slot == 0) {
#pragma unroll
            for (int e = 0; e < EPL; ++e) sm_part[warp][r][li * EPL + e] = o[r][e];
            if (li == 0) { sm_part[warp][r][HD] = m[r]; sm_part[warp][r][HD + 1] = l[r]; }
        }
    }
    __syncthreads();
    // merge the warps of the CTA -> partial record
    const size_t tile_id = (size_t)kvh * row_tiles + rt;
    float* rec = p.ws + (tile_id * p.splits + split) * RT * REC;
    for (int idx = threadIdx.x; idx < RT * HD; idx += blockDim.x) {
        int r = idx / HD, d = idx % HD;
        float mn = kNegBig;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) mn = fmaxf(mn, sm_part[w][r][HD]);
        float ov = 0.f, lv = 0.f;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) {
            float c = exp2f(sm_part[w][r][HD] - mn);
            ov = fmaf(sm_part[w][r][d], c, ov);
            lv = fmaf(sm_part[w][r][HD + 1], c, lv);
        }
        if (p.splits == 1) {
            int row = rt * RT + r;
            if (row < R) {
                int g = row / p.q_len, t = row % p.q_len;
                st_elem<DT>(p.out, (size_t)(kvh * G + g) * p.o_sh + (size_t)t * p.o_st + d, lv > 0.f ? ov / lv : 0.f);
            }
        } else {
            __stcg(rec + r * REC + d, ov);
            if (d == 0) { __stcg(rec + r * REC + HD, mn); __stcg(rec + r * REC + HD + 1, lv); }
        }
    }
    if (p.splits == 1) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_ticket = atomicAdd(p.counters + tile_id, 1u);
    __syncthreads();
    if (s_ticket != (unsigned)(p.splits - 1)) return;
    __threadfence();
    const float* base = p.ws + tile_id * p.splits * RT * REC;
    // last CTA of the tile: fold the partials in split order.  (m, l) of every split are staged in
    // shared memory first so that the fold issues its o-loads 8 at a time instead of one dependent
    // chain of L2 round trips per split.
    __shared__ float s_m[RT][kAttnMaxSplits], s_l[RT][kAttnMaxSplits];
    for (int idx = threadIdx.x; idx < RT * p.splits; idx += blockDim.x) {
        const int r = idx / p.splits, sp = idx % p.splits;
        const float* rr = base + ((size_t)sp * RT + r) * REC;
        s_m[r][sp] = __ldcg(rr + HD);
        s_l[r][sp] = __ldcg(rr + HD + 1);
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < RT * HD; idx += blockDim.x) {
        const int r = idx / HD, d = idx % HD;
        const int row = rt * RT + r;
        if (row >= R) continue;
        float mn = kNegBig;
        for (int sp = 0; sp < p.splits; ++sp) mn = fmaxf(mn, s_m[r][sp]);
        float ov = 0.f, lv = 0.f;
        int sp = 0;
        for (; sp + 8 <= p.splits; sp += 8) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldcg(base + ((size_t)(sp + u) * RT + r) * REC + d);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float c = exp2f(s_m[r][sp + u] - mn);
                ov = fmaf(v[u], c, ov);
                lv = fmaf(s_l[r][sp + u], c, lv);
            }
        }
        for (; sp < p.splits; ++sp) {
            const float c = exp2f(s_m[r][sp] - mn);
            ov = fmaf(__ldcg(base + ((size_t)sp * RT + r) * REC + d), c, ov);
            lv = fmaf(s_l[r][sp], c, lv);
        }
        int g = row / p.q_len, t = row % p.q_len;
        st_elem<DT>(p.out, (size_t)(kvh * G + g) * p.o_sh + (size_t)t * p.o_st + d, lv > 0.f ? ov / lv : 0.f);
    }
    if (threadIdx.x == 0) p.counters[tile_id] = 0;
}

static float* g_attn_ws = nullptr;
static unsigned* g_attn_cnt = nullptr;
static const size_t kAttnWsBytes = 64ull << 20;
static const int kAttnMaxTiles = 16384;

template <int DT, int HD>
static int launch_sdpa(SdpaParams p, cudaStream_t st) {
    constexpr int RT = 4;
    const int G = p.hq / p.n_kv, R = G * p.q_len;
    const int row_tiles = (R + RT - 1) / RT;
    const int base_ctas = p.n_kv * row_tiles;
    // enough CTAs for ~2 per SM, at least 128 positions (two batches of the 8 warps) per split
    int hint = p.context_len > 0 ? p.context_len : p.max_seq;
    int splits = (2 * sm_count() + base_ctas - 1) / base_ctas;
    int max_splits = (hint + 127) / 128;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > kAttnMaxSplits) splits = kAttnMaxSplits;
    PGK_REQUIRE(base_ctas <= kAttnMaxTiles && (size_t)base_ctas * splits * RT * (HD + 2) * 4 <= kAttnWsBytes,
                "sdpa_causal_fixed_cache failed: workspace too small for %d heads x %d rows", p.hq, p.q_len);
    if (!g_attn_ws) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(st, &cs);
        PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                    "sdpa_causal_fixed_cache failed: first call must happen outside graph capture (workspace allocation)");
        PGK_CUDA(cudaMalloc(&g_attn_ws, kAttnWsBytes));
        PGK_CUDA(cudaMalloc(&g_attn_cnt, kAttnMaxTiles * sizeof(unsigned)));
        PGK_CUDA(cudaMemset(g_attn_cnt, 0, kAttnMaxTiles * sizeof(unsigned)));
    }
    p.splits = splits;
    p.ws = g_attn_ws;
    p.counters = g_attn_cnt;
    constexpr size_t kSmem = (size_t)kAttnStages * 2 * AttnTile<DT, HD>::TP * HD * (DT == 0 ? 4 : 2);
    static bool attr_set = false;
    if (!attr_set) {
        PGK_CUDA(cudaFuncSetAttribute(sdpa_decode_kernel<DT, HD, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        attr_set = true;
    }
    PGK_CUDA(launch(sdpa_decode_kernel<DT, HD, RT>, dim3(splits, row_tiles, p.n_kv), dim3(kAttnWarps * 32), kSmem, st, p));
    return 0;
}


// ---------------------------------------------------------------------------------------------------
// Fused decode bracket + attention for the q_len == 1, bf16, head_dim 128 case (every config of
// BASELINE.json): [QK-RMSNorm] -> RoPE -> K/V cache write -> split-KV attention in ONE launch.
// Replaces the chain rmsnorm(q), rmsnorm(k), rope_inplace_f32table, kv_cache_update_gqa x2,
// sdpa_causal_fixed_cache_ptr of layers/attention.py:326-368 with the same rounding points (q/k rounded
// to bf16 after the norm and after the rotation, probabilities and the output accumulator in f32).
//   * K/V tiles of the OLD positions do not depend on the previous kernel (the qkv GEMV), so one thread
//     reads the device-side context length and issues the TMA loads of the CTA's first stages BEFORE
//     griddepcontrol.wait: the HBM round trip overlaps the GEMV's tail.
//   * after the wait the consumer warps normalise + rotate the query heads of the tile into shared memory, the
//     producer warp prepares the new K / V row; the warp whose positions hold the new one patches the row
//     into its staged tile (and the rt == 0 CTA writes it to the cache) -- no kernel boundary between the
//     cache write and the attention that reads it.
struct RopeSdpaParams {
    const uint16_t* qkv;                       // [(hq + 2 n_kv) * 128] raw projections of the new token
    const uint16_t* q_gamma; const uint16_t* k_gamma; float eps;
    const float* cos_t; const float* sin_t;    // [positions, 128] f32
    uint16_t* kc; uint16_t* vc;                // [n_kv, max_seq, 128]
    uint16_t* out;                             // [hq * 128]
    int hq, n_kv, max_seq, pos, context_len;
    const int32_t* pos_buf; const int32_t* ctx_buf;
    float scale_log2;
    int splits;
    uint2* ll; unsigned* gen;                  // LL partial words {f32 bits, exchange number}, per-record generation
    int box_rows;                              // rows of one TMA box = min(64, max_seq)
    // multi-sequence decode (continuous batching, native/ops/batch/continuous_batching.cuh:70-245): n_seq independent
    // sequences, each with its own cache region, position and context length (pos_buf / ctx_buf are then arrays)
    int n_seq;
    long qkv_stride, out_stride, cache_stride;     // elements between consecutive sequences
};
// ---------------------------------------------------------------------------------------------------------------
// The score and P.V products run on the tensor cores (mma.sync m16n8k16, SASS HMMA)
// and the K/V tiles fetched by tensor-map TMA (cp.async.bulk.tensor.4d, SWIZZLE_128B, SASS UTMALDG) so that ldmatrix
// reads them without bank conflicts.  Decode attention is HBM / latency bound; the CUDA-core loop of round 1 spent 450
// issue slots per warp per 64 positions (unpacking V, 15 shuffles of the score reduce-scatter), which bounded the long
// contexts and the multi-sequence batches.  Here a warp owns 16 positions of a 64-position tile:
//     S[8 heads (+8 zero rows), 16 pos]  = Q[16, 128] . K^T      16 HMMA, B fragments by ldmatrix from the swizzled tile
//     online softmax on the 4 scores a thread holds (row = lane / 4), quad shuffles only
//     O[16, 128] += P[16, 16] . V[16, 128]                       16 HMMA, V fragments by ldmatrix.trans
// The probabilities go to the tensor cores as TWO bf16 terms, p = hi + lo, hi in rows 0..7 and lo in rows 8..15 of the
// A fragment (the rows the <= 8 query heads leave empty), so one MMA accumulates both and O = O[0..7] + O[8..15] carries
// 16 bits of p: the products q.k are exact and p.v is as accurate as the f32 loop it replaces.
// One CTA = 4 consumer warps + 1 TMA producer warp, 2 CTAs per SM; up to 8 query heads per CTA, so a GQA group of 8
// (Llama-3-70B, Qwen3-30B-A3B) reads its K/V once instead of twice.  Stages: 2 x 32 KB for one sequence -- measured:
// with 3 stages the two CTAs need 202 KB of shared memory, the SM is switched to its largest carve-out, and the GEMVs
// that FOLLOW in the step run 8 % slower on the 28 KB of L1 that leaves them (profiles/r02_attn_mma_carveout.txt) --
// and 3 x 32 KB for multi-sequence batches, where attention is most of the step.
constexpr int kMW = 4;                          // consumer warps
constexpr int kMTile = 64;                      // KV positions per stage
constexpr int kMRT = 8;                         // query heads per CTA
constexpr int kMThreads = (kMW + 1) * 32;
constexpr uint32_t kMBox = 64 * 128;            // one TMA box: 64 positions x 64 dims (128 bytes, the swizzle span)
constexpr uint32_t kMStageB = 4 * kMBox;        // K dims 0..63 | K dims 64..127 | V dims 0..63 | V dims 64..127
constexpr int kMRec = 130;                      // partial record: o[128], m, l

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void tma_load_4d(uint32_t smem_dst, const CUtensorMap* map, int c0, int c1, int c2, int c3, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                 :: "r"(smem_dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ll_store(uint2* p, uint32_t data, uint32_t seq) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(p), "r"(data), "r"(seq) : "memory");
}
// a wait that cannot hang the device: a protocol bug traps instead of spinning forever
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (ok) return;
    }
    __trap();
}
// [RMSNorm] + NeoX rotation of one 128-wide head row already in registers (x[i] = row[lane + 32 i]); same arithmetic
// and rounding points as norm_rope_head128 above
__device__ __forceinline__ void norm_rope_regs(float (&x)[4], bool use_gamma, const float (&g)[4], float eps, const float (&c)[2],
                                               const float (&s)[2], int lane, uint16_t* __restrict__ dst_s, uint16_t* __restrict__ dst_g) {
    if (use_gamma) {
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            float part = warp_sum(fmaf(x[w], x[w], 0.f));
            if (lane == w) tot = part;
        }
        const float inv = rsqrtf(warp_sum(tot) / 128.f + eps);
#pragma unroll
        for (int i = 0; i < 4; ++i) x[i] = round_bf16(x[i] * inv * g[i]);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float x0 = x[i], x1 = x[i + 2];
        const uint16_t y0 = f32_to_bf16_bits(x0 * c[i] - x1 * s[i]), y1 = f32_to_bf16_bits(x1 * c[i] + x0 * s[i]);
        dst_s[lane + 32 * i] = y0; dst_s[lane + 32 * i + 64] = y1;
        if (dst_g) { dst_g[lane + 32 * i] = y0; dst_g[lane + 32 * i + 64] = y1; }
    }
}

template <int kMStages>
__global__ void __launch_bounds__(kMThreads, 2)
rope_sdpa_mma_kernel(const __grid_constant__ CUtensorMap kmap, const __grid_constant__ CUtensorMap vmap, RopeSdpaParams p) {
    constexpr int HD = 128, RT = kMRT, REC = kMRec;
    constexpr int SREC = 136;                                  // row stride of the CTA merge buffer: float2 stores of a half-warp stay conflict-free
    extern __shared__ unsigned char dyn_raw[];
    __shared__ __align__(16) uint16_t q_s[RT][HD + 8];          // +16 bytes per row: the fragment loads of 8 rows x 4 words hit 32 banks
    __shared__ __align__(16) uint16_t new_s[2][HD];           // new K row, new V row
    __shared__ __align__(8) uint64_t full_bar[kMStages];
    __shared__ __align__(8) uint64_t empty_bar[kMStages];
    __shared__ unsigned s_gen[kAttnMaxSplits];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = p.hq / p.n_kv;
    const int row_tiles = (G + RT - 1) / RT;
    const int split = blockIdx.x, rt = blockIdx.y, kvh = blockIdx.z % p.n_kv, seq = blockIdx.z / p.n_kv;
    const int rows = min(RT, G - rt * RT);                    // query heads this CTA serves
    p.qkv += (size_t)seq * p.qkv_stride;
    p.out += (size_t)seq * p.out_stride;
    p.kc += (size_t)seq * p.cache_stride;
    p.vc += (size_t)seq * p.cache_stride;
    const uint32_t tiles_s = (smem_u32(dyn_raw) + 1023u) & ~1023u;      // the swizzle pattern repeats every 1024 bytes
    unsigned char* tiles_g = dyn_raw + (tiles_s - smem_u32(dyn_raw));

    // ---- independent of the previous kernel: position, barriers, the first K/V tiles, tables
    const int pos = p.pos_buf ? p.pos_buf[seq] : p.pos;
    const int ctx = min(p.ctx_buf ? p.ctx_buf[seq] : p.context_len, p.max_seq);
    int chunk = (ctx + p.splits - 1) / p.splits;
    chunk = (chunk + 15) & ~15;
    const int c0 = split * chunk, c1 = min(ctx, c0 + chunk);
    const int ntiles = (c1 > c0) ? (c1 - c0 + kMTile - 1) / kMTile : 0;
    const bool has_new = pos >= c0 && pos < c1;
    const bool writer = rt == 0 && pos < p.max_seq && (has_new || (pos >= ctx && split == 0));
    auto issue_tile = [&](int t) {
        const int st = t % kMStages;
        const int t0 = c0 + t * kMTile;
        const uint32_t dst = tiles_s + (uint32_t)st * kMStageB;
        mbar_arrive_expect_tx(&full_bar[st], (uint32_t)p.box_rows * 512u);   // 4 boxes of box_rows x 128 bytes; out-of-range rows are zero-filled and still counted
        tma_load_4d(dst, &kmap, 0, t0, kvh, seq, &full_bar[st]);
        tma_load_4d(dst + kMBox, &kmap, 64, t0, kvh, seq, &full_bar[st]);
        tma_load_4d(dst + 2 * kMBox, &vmap, 0, t0, kvh, seq, &full_bar[st]);
        tma_load_4d(dst + 3 * kMBox, &vmap, 64, t0, kvh, seq, &full_bar[st]);
    };
    if (threadIdx.x == kMW * 32) {
#pragma unroll
        for (int i = 0; i < kMStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], kMW); }
        mbar_fence_init();
        asm volatile("prefetch.tensormap [%0];" :: "l"(&kmap) : "memory");
        asm volatile("prefetch.tensormap [%0];" :: "l"(&vmap) : "memory");
        for (int t = 0; t < min(ntiles, kMStages); ++t) issue_tile(t);
    }
    float rc[2], rs[2], rg[4] = {1.f, 1.f, 1.f, 1.f};
    const uint16_t* my_gamma = (warp < kMW) ? p.q_gamma : p.k_gamma;
    {
        const float* cs = p.cos_t + (size_t)pos * HD;
        const float* sn = p.sin_t + (size_t)pos * HD;
#pragma unroll
        for (int i = 0; i < 2; ++i) { rc[i] = __ldg(cs + lane + 32 * i); rs[i] = __ldg(sn + lane + 32 * i); }
        if (my_gamma) {
#pragma unroll
            for (int i = 0; i < 4; ++i) rg[i] = bf16_bits_to_f32(__ldg(my_gamma + lane + 32 * i));
        }
    }
    griddep_wait();
    griddep_launch();
    // exchange numbers of this tile's partial records (see the fold below)
    if (p.splits > 1 && threadIdx.x < p.splits)
        s_gen[threadIdx.x] = __ldcg(p.gen + (((size_t)seq * p.n_kv + kvh) * row_tiles + rt) * p.splits + threadIdx.x) + 1u;

    // ---- the new token: consumer warp w rotates query heads w and w + 4, the producer warp the new K / V rows
    if (warp < kMW) {
        float x[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int h = warp + j * kMW;
            const uint16_t* row = p.qkv + (size_t)(kvh * G + rt * RT + min(h, rows - 1)) * HD;
#pragma unroll
            for (int i = 0; i < 4; ++i) x[j][i] = bf16_bits_to_f32(row[lane + 32 * i]);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int h = warp + j * kMW;
            if (h < rows) norm_rope_regs(x[j], my_gamma != nullptr, rg, p.eps, rc, rs, lane, q_s[h], nullptr);
            else for (int d = lane; d < HD; d += 32) q_s[h][d] = 0;
        }
    } else if (has_new || writer) {
        float x[4];
        const uint16_t* krow = p.qkv + (size_t)(p.hq + kvh) * HD;
        const uint16_t* vrow = p.qkv + (size_t)(p.hq + p.n_kv + kvh) * HD;
#pragma unroll
        for (int i = 0; i < 4; ++i) x[i] = bf16_bits_to_f32(krow[lane + 32 * i]);
        uint16_t vv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) vv[i] = vrow[lane + 32 * i];
        uint16_t* kdst = writer ? p.kc + ((size_t)kvh * p.max_seq + pos) * HD : nullptr;
        uint16_t* vdst = writer ? p.vc + ((size_t)kvh * p.max_seq + pos) * HD : nullptr;
        norm_rope_regs(x, my_gamma != nullptr, rg, p.eps, rc, rs, lane, new_s[0], kdst);
#pragma unroll
        for (int i = 0; i < 4; ++i) { new_s[1][lane + 32 * i] = vv[i]; if (vdst) vdst[lane + 32 * i] = vv[i]; }
    }
    __syncthreads();   // q_s / new_s visible; also orders the mbarrier init before every wait

    float o[16][4];                                            // [dim tile of 8][c0 c1: hi rows | c2 c3: lo rows]
    float m_run = kNegBig, l_run = 0.f;
    if (warp == kMW) {
        // ---- producer: refill a stage as soon as the four consumer warps have released it
        if (lane == 0) {
            for (int t = kMStages; t < ntiles; ++t) {
                mbar_wait_bounded(&empty_bar[t % kMStages], (uint32_t)((t / kMStages - 1) & 1));
                issue_tile(t);
            }
        }
    } else {
        const int qrow = lane >> 2, kk = (lane & 3) * 2;
        uint32_t qa[8][2];
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            qa[ks][0] = *reinterpret_cast<const uint32_t*>(&q_s[qrow][ks * 16 + kk]);
            qa[ks][1] = *reinterpret_cast<const uint32_t*>(&q_s[qrow][ks * 16 + kk + 8]);
        }
#pragma unroll
        for (int nt = 0; nt < 16; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) o[nt][e] = 0.f;
        // ldmatrix row addresses of this lane (matrix j = lane / 8, row i = lane % 8), relative to the tile
        const int lj = lane >> 3, lr8 = lane & 7;
        const int k_row = warp * 16 + (lj >> 1) * 8 + lr8;      // K: matrices (n-tile 0, k lo) (n-tile 0, k hi) (n-tile 1, k lo) (n-tile 1, k hi)
        const int v_row = warp * 16 + (lj & 1) * 8 + lr8;       // V: (pos 0..7, dims) (pos 8..15, dims) (pos 0..7, dims + 8) (pos 8..15, dims + 8)
        const uint32_t k_off = (uint32_t)k_row * 128u, v_off = 2 * kMBox + (uint32_t)v_row * 128u;
        const uint32_t k_sw = (uint32_t)(k_row & 7), v_sw = (uint32_t)(v_row & 7);

        for (int t = 0; t < ntiles; ++t) {
            const int st = t % kMStages;
            const int t0 = c0 + t * kMTile;
            const int wbase = t0 + warp * 16;
            mbar_wait_bounded(&full_bar[st], (uint32_t)((t / kMStages) & 1));
            const uint32_t tile = tiles_s + (uint32_t)st * kMStageB;
            unsigned char* tile_g = tiles_g + (size_t)st * kMStageB;
            bool wrote = false;
            if (wbase + 16 > ctx && wbase < c1) {               // warp-uniform: V rows past the context may hold anything
                for (int i = lane; i < 16 * 16; i += 32) {
                    const int r = warp * 16 + (i >> 4), ch = i & 15;
                    if (t0 + r >= ctx)
                        *reinterpret_cast<uint4*>(tile_g + 2 * kMBox + (ch >> 3) * kMBox + r * 128 + (ch & 7) * 16) = make_uint4(0, 0, 0, 0);
                }
                wrote = true;
            }
            if (has_new && pos >= wbase && pos < wbase + 16) {  // warp-uniform: patch the new row into the staged tile
                const int r = pos - t0, which = lane >> 4, ch = lane & 15;
                *reinterpret_cast<uint4*>(tile_g + which * 2 * kMBox + (ch >> 3) * kMBox + r * 128 + (((ch & 7) ^ (r & 7)) * 16)) =
                    reinterpret_cast<const uint4*>(new_s[which])[ch];
                wrote = true;
            }
            if (wrote) __syncwarp();
            if (wbase < c1) {
                // ---- scores of 16 positions x 8 heads
                float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    const uint32_t ch = (uint32_t)(ks * 2 + (lj & 1));
                    uint32_t b[4];
                    ldsm_x4(tile + (ch >> 3) * kMBox + k_off + (((ch & 7) ^ k_sw) << 4), b);
                    mma_bf16_16816(s0, qa[ks][0], 0u, qa[ks][1], 0u, b[0], b[1]);
                    mma_bf16_16816(s1, qa[ks][0], 0u, qa[ks][1], 0u, b[2], b[3]);
                }
                const int pj = wbase + kk;                       // positions pj, pj + 1 (s0) and pj + 8, pj + 9 (s1)
                float sc[4];
                sc[0] = pj < c1 ? s0[0] * p.scale_log2 : kNegBig;
                sc[1] = pj + 1 < c1 ? s0[1] * p.scale_log2 : kNegBig;
                sc[2] = pj + 8 < c1 ? s1[0] * p.scale_log2 : kNegBig;
                sc[3] = pj + 9 < c1 ? s1[1] * p.scale_log2 : kNegBig;
                float mx = fmaxf(fmaxf(sc[0], sc[1]), fmaxf(sc[2], sc[3]));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                const float mn = fmaxf(m_run, mx);
                const float corr = exp2f(m_run - mn);
                float pr[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) pr[j] = (sc[j] > 0.5f * kNegBig) ? exp2f(sc[j] - mn) : 0.f;
                l_run = l_run * corr + ((pr[0] + pr[1]) + (pr[2] + pr[3]));
                m_run = mn;
                if (__any_sync(0xffffffffu, corr != 1.f)) {
#pragma unroll
                    for (int nt = 0; nt < 16; ++nt)
#pragma unroll
                        for (int e = 0; e < 4; ++e) o[nt][e] *= corr;
                }
                // p = hi + lo as two bf16 terms: hi in rows 0..7, lo in rows 8..15 of the A fragment
                float lo[4];
                uint32_t hb[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) { hb[j] = f32_to_bf16_bits(pr[j]); lo[j] = pr[j] - bf16_bits_to_f32(hb[j]); }
                const uint32_t a0 = hb[0] | (hb[1] << 16), a2 = hb[2] | (hb[3] << 16);
                const uint32_t a1 = pack_bf16x2(lo[0], lo[1]), a3 = pack_bf16x2(lo[2], lo[3]);
                // ---- O += P . V
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const uint32_t ch = (uint32_t)(q * 2 + (lj >> 1));
                    uint32_t b[4];
                    ldsm_x4_t(tile + (ch >> 3) * kMBox + v_off + (((ch & 7) ^ v_sw) << 4), b);
                    mma_bf16_16816(o[2 * q], a0, a1, a2, a3, b[0], b[1]);
                    mma_bf16_16816(o[2 * q + 1], a0, a1, a2, a3, b[2], b[3]);
                }
            }
            if (t + kMStages < ntiles) {                        // hand the stage back to the producer
                if (wrote) fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[st]);
            }
        }
        l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
        l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    }
    __syncthreads();                                           // every stage is consumed: its memory is free for the merge
    float (*sm_part)[RT][SREC] = reinterpret_cast<float (*)[RT][SREC]>(tiles_g);
    if (warp < kMW) {
        const int qrow = lane >> 2, kk = (lane & 3) * 2;
#pragma unroll
        for (int nt = 0; nt < 16; ++nt)
            *reinterpret_cast<float2*>(&sm_part[warp][qrow][nt * 8 + kk]) = make_float2(o[nt][0] + o[nt][2], o[nt][1] + o[nt][3]);
        if ((lane & 3) == 0) { sm_part[warp][qrow][HD] = m_run; sm_part[warp][qrow][HD + 1] = l_run; }
    }
    __syncthreads();
    const size_t tile_id = ((size_t)seq * p.n_kv + kvh) * row_tiles + rt;
    if (p.splits == 1) {
        for (int idx = threadIdx.x; idx < rows * HD; idx += kMThreads) {
            const int r = idx / HD, d = idx % HD;
            float mn = kNegBig;
#pragma unroll
            for (int w = 0; w < kMW; ++w) mn = fmaxf(mn, sm_part[w][r][HD]);
            float ov = 0.f, lv = 0.f;
#pragma unroll
            for (int w = 0; w < kMW; ++w) {
                const float c = exp2f(sm_part[w][r][HD] - mn);
                ov = fmaf(sm_part[w][r][d], c, ov);
                lv = fmaf(sm_part[w][r][HD + 1], c, lv);
            }
            p.out[(size_t)(kvh * G + rt * RT + r) * HD + d] = f32_to_bf16_bits(lv > 0.f ? ov / lv : 0.f);
        }
        return;
    }
    // ---- cross-split fold without a fence, a ticket or a second pass: every partial travels as 8-byte words
    // {f32 bits, exchange number} (the LL form of p2p.cu), so a reader knows a word is this launch's by looking at it.
    // Each partial record (tile, split) of the workspace has its own generation counter; the exchange number is that
    // generation + 1, read by every CTA of the tile after the dependency wait.  CTA `split` folds the slice
    // [split * W, (split + 1) * W) of the rows * 128 outputs in split order (all slices poll at once: one L2 round
    // trip after the slowest writer); split 0 publishes the new generations once its slice has seen every partial,
    // which proves that every CTA of the tile has read the old ones.  The numbers written to any one address only
    // ever grow, whatever shapes earlier launches had, so a stale word can never pass for a fresh one.
    // (A fold CTA waits for its sibling CTAs: one tile's splits <= 128 are always co-resident, blocks being
    // dispatched in index order; the wait is bounded and traps.)
    const int S = p.splits;
    uint2* ll = p.ll + tile_id * (size_t)S * RT * REC;
    for (int idx = threadIdx.x; idx < rows * HD; idx += kMThreads) {
        const int r = idx / HD, d = idx % HD;
        float mn = kNegBig;
#pragma unroll
        for (int w = 0; w < kMW; ++w) mn = fmaxf(mn, sm_part[w][r][HD]);
        float ov = 0.f, lv = 0.f;
#pragma unroll
        for (int w = 0; w < kMW; ++w) {
            const float c = exp2f(sm_part[w][r][HD] - mn);
            ov = fmaf(sm_part[w][r][d], c, ov);
            lv = fmaf(sm_part[w][r][HD + 1], c, lv);
        }
        uint2* rec = ll + ((size_t)split * RT + r) * REC;
        const unsigned tag = s_gen[split];
        ll_store(rec + d, __float_as_uint(ov), tag);
        if (d == 0) { ll_store(rec + HD, __float_as_uint(mn), tag); ll_store(rec + HD + 1, __float_as_uint(lv), tag); }
    }
    const int E = rows * HD;
    const int F = S;
    const int W = (E + F - 1) / F;
    const int e0 = split * W, n_e = min(E, e0 + W) - e0;
    if (n_e <= 0) return;
    const int r0 = e0 / HD, nr = (e0 + n_e - 1) / HD - r0 + 1;   // rows this slice touches
    const int SP = S | 1;                                        // odd stride: conflict-free column walks
    float* f_o = reinterpret_cast<float*>(tiles_g + 20 * 1024);  // [n_e][SP]
    float* f_m = f_o + W * SP;                                   // [RT][SP]  (max, then scale factor)
    float* f_l = f_m + RT * SP;                                  // [RT][SP]
    float* f_den = f_l + RT * SP;                                // [RT]
    constexpr int MAXJ = 16;
    const int n_o = n_e * S, n_items = n_o + nr * S * 2;
    for (int base = 0; base < n_items; base += kMThreads * MAXJ) {
        uint32_t src[MAXJ], tag[MAXJ];
        int dst[MAXJ];                                           // >= 0: f_o index; < 0: ~index into f_m (.. f_l follows)
        unsigned pend = 0;
#pragma unroll
        for (int j = 0; j < MAXJ; ++j) {
            const int k = base + threadIdx.x + j * kMThreads;
            src[j] = 0; dst[j] = 0; tag[j] = 0;
            if (k < n_o) {
                const int el = k / S, sp = k - el * S, e = e0 + el;
                src[j] = (uint32_t)((sp * RT + (e >> 7)) * REC + (e & 127));
                dst[j] = el * SP + sp;
                tag[j] = s_gen[sp];
                pend |= 1u << j;
            } else if (k < n_items) {
                const int k2 = k - n_o, which = k2 & 1, k3 = k2 >> 1;
                const int rr = k3 / S, sp = k3 - rr * S;
                src[j] = (uint32_t)((sp * RT + r0 + rr) * REC + HD + which);
                dst[j] = ~(which * RT * SP + (r0 + rr) * SP + sp);
                tag[j] = s_gen[sp];
                pend |= 1u << j;
            }
        }
        for (uint32_t spin = 0; pend; ++spin) {
            uint2 w[MAXJ];
#pragma unroll
            for (int j = 0; j < MAXJ; ++j)
                if (pend >> j & 1)
                    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(w[j].x), "=r"(w[j].y) : "l"(ll + src[j]) : "memory");
#pragma unroll
            for (int j = 0; j < MAXJ; ++j)
                if ((pend >> j & 1) && w[j].y == tag[j]) {
                    if (dst[j] >= 0) f_o[dst[j]] = __uint_as_float(w[j].x); else f_m[~dst[j]] = __uint_as_float(w[j].x);
                    pend &= ~(1u << j);
                }
            if (spin > (1u << 22)) __trap();                     // a lost partial must not hang the device
        }
    }
    __syncthreads();
    if (split == 0 && threadIdx.x < S) p.gen[tile_id * S + threadIdx.x] = s_gen[threadIdx.x];
    // per row: max over the splits, scale factors, denominator (one warp per row, fixed shuffle tree)
    for (int rr = warp; rr < nr; rr += kMW + 1) {
        const int r = r0 + rr;
        float mx = kNegBig;
        for (int sp = lane; sp < S; sp += 32) mx = fmaxf(mx, f_m[r * SP + sp]);
        mx = warp_max(mx);
        float den = 0.f;
        for (int sp = lane; sp < S; sp += 32) {
            const float c = exp2f(f_m[r * SP + sp] - mx);
            f_m[r * SP + sp] = c;
            den = fmaf(f_l[r * SP + sp], c, den);
        }
        den = warp_sum(den);
        if (lane == 0) f_den[r] = den;
    }
    __syncthreads();
    for (int el = threadIdx.x; el < n_e; el += kMThreads) {
        const int e = e0 + el, r = e >> 7, d = e & 127;
        const float* fo = f_o + el * SP;
        const float* fc = f_m + r * SP;
        float ov = 0.f;
        for (int sp = 0; sp < S; ++sp) ov = fmaf(fo[sp], fc[sp], ov);
        const float lv = f_den[r];
        p.out[(size_t)(kvh * G + rt * RT + r) * HD + d] = f32_to_bf16_bits(lv > 0.f ? ov / lv : 0.f);
    }
}

// ---- host side of the tensor-map path
static uint2* g_attn_ll = nullptr;
static unsigned* g_attn_gen = nullptr;
static const size_t kAttnLLBytes = 64ull << 20;
static const size_t kAttnLLRecords = kAttnLLBytes / (kMRT * kMRec * 8);     // partial records (tile, split) the workspace holds
typedef CUresult (*AttnEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                      const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                      CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static AttnEncodeTiledFn attn_encode() {
    static AttnEncodeTiledFn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return (AttnEncodeTiledFn)f;
    }();
    return fn;
}
// [n_seq, n_kv, max_seq, 128] bf16 cache as a 4-D tensor, box = 64 dims x min(64, max_seq) positions, 128-byte swizzle
// (cached per cache)
static int attn_cache_map(const void* cache, int max_seq, int n_kv, int n_seq, long cache_stride, CUtensorMap* out) {
    struct Key { const void* p; int a, b, c; long s; };
    static std::mutex mu;
    static std::vector<std::pair<Key, CUtensorMap>> maps;
    std::lock_guard<std::mutex> lk(mu);
    for (auto& e : maps)
        if (e.first.p == cache && e.first.a == max_seq && e.first.b == n_kv && e.first.c == n_seq && e.first.s == cache_stride) { *out = e.second; return 0; }
    CUtensorMap m;
    const cuuint64_t dims[4] = {128, (cuuint64_t)max_seq, (cuuint64_t)n_kv, (cuuint64_t)n_seq};
    const cuuint64_t strides[3] = {256, (cuuint64_t)max_seq * 256, (cuuint64_t)(n_seq > 1 ? cache_stride * 2 : (long)n_kv * max_seq * 256)};
    const cuuint32_t box[4] = {64, (cuuint32_t)(max_seq < kMTile ? max_seq : kMTile), 1, 1}, estr[4] = {1, 1, 1, 1};
    const CUresult r = attn_encode()(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(cache), dims, strides, box, estr,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    PGK_REQUIRE(r == CUDA_SUCCESS, "rope_sdpa_decode failed: cuTensorMapEncodeTiled returned %d", (int)r);
    if (maps.size() >= 4096) maps.clear();
    maps.emplace_back(Key{cache, max_seq, n_kv, n_seq, cache_stride}, m);
    *out = m;
    return 0;
}

template <int STAGES>
static int launch_rope_sdpa_stages(const CUtensorMap& kmap, const CUtensorMap& vmap, const RopeSdpaParams& p, dim3 grid, cudaStream_t st) {
    constexpr size_t kSmem = (size_t)STAGES * kMStageB + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        PGK_CUDA(cudaFuncSetAttribute(rope_sdpa_mma_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        attr_set = true;
    }
    PGK_CUDA(launch(rope_sdpa_mma_kernel<STAGES>, grid, dim3(kMThreads), kSmem, st, kmap, vmap, p));
    return 0;
}

static int launch_rope_sdpa(RopeSdpaParams p, cudaStream_t st) {
    PGK_REQUIRE(attn_encode() != nullptr, "rope_sdpa_decode failed: the driver does not export cuTensorMapEncodeTiled");
    const int G = p.hq / p.n_kv;
    const int row_tiles = (G + kMRT - 1) / kMRT;
    if (p.n_seq < 1) p.n_seq = 1;
    const int base_ctas = p.n_kv * row_tiles * p.n_seq;
    // two resident CTAs per SM; a split keeps at least one 64-position tile
    static const int env_splits = [] { const char* e = getenv("PGK_ATTN_SPLITS"); return e ? atoi(e) : 0; }();
    int splits = (2 * sm_count() + base_ctas - 1) / base_ctas;
    const int hint = p.ctx_buf ? p.max_seq : p.context_len;
    const int max_splits = (hint + kMTile - 1) / kMTile;
    if (splits > max_splits) splits = max_splits;
    if (splits > 32) splits = 32;           // measured: beyond 32 the fold grows faster than the tiles shrink (A3B 4 KV heads: 64 -> 32 splits +1.5 %)
    if (env_splits > 0) splits = env_splits;
    if (splits < 1) splits = 1;
    if (splits > kAttnMaxSplits) splits = kAttnMaxSplits;
    PGK_REQUIRE(base_ctas <= kAttnMaxTiles && (size_t)base_ctas * splits <= kAttnLLRecords,
                "rope_sdpa_decode failed: too many (sequence, head) tiles (%d)", base_ctas);
    if (!g_attn_ll) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(st, &cs);
        PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                    "rope_sdpa_decode failed: first call must happen outside graph capture (workspace allocation)");
        PGK_CUDA(cudaMalloc(&g_attn_ll, kAttnLLBytes));
        PGK_CUDA(cudaMemset(g_attn_ll, 0, kAttnLLBytes));
        PGK_CUDA(cudaMalloc(&g_attn_gen, kAttnLLRecords * sizeof(unsigned)));
        PGK_CUDA(cudaMemset(g_attn_gen, 0, kAttnLLRecords * sizeof(unsigned)));
    }
    CUtensorMap kmap, vmap;
    if (attn_cache_map(p.kc, p.max_seq, p.n_kv, p.n_seq, p.cache_stride, &kmap)) return 1;
    if (attn_cache_map(p.vc, p.max_seq, p.n_kv, p.n_seq, p.cache_stride, &vmap)) return 1;
    p.splits = splits;
    p.ll = g_attn_ll;
    p.gen = g_attn_gen;
    p.box_rows = p.max_seq < kMTile ? p.max_seq : kMTile;
    const dim3 grid(splits, row_tiles, p.n_kv * p.n_seq);
    return p.n_seq > 1 ? launch_rope_sdpa_stages<3>(kmap, vmap, p, grid, st) : launch_rope_sdpa_stages<2>(kmap, vmap, p, grid, st);
}

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_sdpa_decode(const void* Q, const void* K, const void* V, void* out, int n_heads_q, int n_kv, int q_len,
                    int head_dim, int max_seq, int context_len, const int32_t* ctx_buf, float scale,
                    long q_stride_h, long q_stride_t, long o_stride_h, long o_stride_t, int dtype, void* stream) {
    PGK_REQUIRE(n_kv > 0 && n_heads_q % n_kv == 0, "sdpa_causal_fixed_cache failed: n_heads %d not a multiple of n_kv_heads %d", n_heads_q, n_kv);
    PGK_REQUIRE(q_len >= 1, "sdpa_causal_fixed_cache failed: q_len must be >= 1");
    PGK_REQUIRE(ctx_buf || (context_len >= 1 && context_len <= max_seq),
                "sdpa_causal_fixed_cache failed: context_len %d out of range (max_seq_len %d)", context_len, max_seq);
    PGK_REQUIRE(((uintptr_t)K % 16 == 0) && ((uintptr_t)V % 16 == 0), "sdpa_causal_fixed_cache failed: K/V must be 16-byte aligned");
    if (scale <= 0.f) scale = 1.0f / sqrtf((float)head_dim);   // sdpa_causal.inl:464-467
    SdpaParams p = {};
    p.Q = Q; p.K = K; p.V = V; p.out = out;
    p.hq = n_heads_q; p.n_kv = n_kv; p.q_len = q_len; p.max_seq = max_seq; p.context_len = context_len;
    p.ctx_buf = ctx_buf;
    p.scale_log2 = scale * 1.4426950408889634f;
    p.q_sh = q_stride_h; p.q_st = q_stride_t; p.o_sh = o_stride_h; p.o_st = o_stride_t;
    cudaStream_t st = (cudaStream_t)stream;
#define PGK_SDPA(DTV, HDV) if (dtype == DTV && head_dim == HDV) return launch_sdpa<DTV, HDV>(p, st);
    PGK_SDPA(2, 128) PGK_SDPA(2, 64) PGK_SDPA(1, 128) PGK_SDPA(1, 64) PGK_SDPA(0, 128) PGK_SDPA(0, 64)
    PGK_SDPA(2, 256) PGK_SDPA(1, 256) PGK_SDPA(2, 32) PGK_SDPA(1, 32) PGK_SDPA(0, 32)
#undef PGK_SDPA
    set_error("sdpa_causal_fixed_cache failed: unsupported dtype %d / head_dim %d", dtype, head_dim);
    return 1;
}

int pgk_rope_sdpa_decode(const void* qkv, const void* q_gamma, const void* k_gamma, float eps, const float* cos_table,
                         const float* sin_table, void* k_cache, void* v_cache, void* out, int n_heads_q, int n_kv,
                         int head_dim, int max_seq, int pos, const int32_t* pos_buf, int context_len,
                         const int32_t* ctx_buf, float scale, void* stream) {
    PGK_REQUIRE(head_dim == 128, "rope_sdpa_decode failed: head_dim %d (the fused bracket is built for 128)", head_dim);
    PGK_REQUIRE(n_kv > 0 && n_heads_q % n_kv == 0, "rope_sdpa_decode failed: n_heads %d not a multiple of n_kv_heads %d", n_heads_q, n_kv);
    PGK_REQUIRE(ctx_buf || (context_len >= 1 && context_len <= max_seq),
                "rope_sdpa_decode failed: context_len %d out of range (max_seq_len %d)", context_len, max_seq);
    PGK_REQUIRE(pos_buf || (pos >= 0 && pos < max_seq), "rope_sdpa_decode failed: position %d out of range", pos);
    PGK_REQUIRE(((uintptr_t)k_cache % 16 == 0) && ((uintptr_t)v_cache % 16 == 0) && ((uintptr_t)qkv % 16 == 0),
                "rope_sdpa_decode failed: qkv / K / V must be 16-byte aligned");
    if (scale <= 0.f) scale = 1.0f / sqrtf((float)head_dim);
    RopeSdpaParams p = {};
    p.qkv = (const uint16_t*)qkv; p.q_gamma = (const uint16_t*)q_gamma; p.k_gamma = (const uint16_t*)k_gamma; p.eps = eps;
    p.cos_t = cos_table; p.sin_t = sin_table; p.kc = (uint16_t*)k_cache; p.vc = (uint16_t*)v_cache; p.out = (uint16_t*)out;
    p.hq = n_heads_q; p.n_kv = n_kv; p.max_seq = max_seq; p.pos = pos; p.context_len = context_len;
    p.pos_buf = pos_buf; p.ctx_buf = ctx_buf;
    p.scale_log2 = scale * 1.4426950408889634f;
    return launch_rope_sdpa(p, (cudaStream_t)stream);
}

/* n_seq independent sequences in one launch (continuous batching): row s of qkv / out, cache region s (cache_stride
 * elements apart: [n_seq, n_kv, max_seq, 128]), pos_buf[s], ctx_buf[s] */
int pgk_rope_sdpa_decode_batch(const void* qkv, const void* q_gamma, const void* k_gamma, float eps, const float* cos_table,
                               const float* sin_table, void* k_cache, void* v_cache, void* out, int n_heads_q, int n_kv,
                               int head_dim, int max_seq, int n_seq, const int32_t* pos_buf, const int32_t* ctx_buf,
                               long cache_stride, float scale, void* stream) {
    PGK_REQUIRE(head_dim == 128, "rope_sdpa_decode_batch failed: head_dim %d (the fused bracket is built for 128)", head_dim);
    PGK_REQUIRE(n_kv > 0 && n_heads_q % n_kv == 0, "rope_sdpa_decode_batch failed: n_heads %d not a multiple of n_kv_heads %d", n_heads_q, n_kv);
    PGK_REQUIRE(n_seq >= 1 && pos_buf && ctx_buf, "rope_sdpa_decode_batch failed: needs n_seq >= 1 and per-sequence position / context arrays");
    PGK_REQUIRE(((uintptr_t)k_cache % 16 == 0) && ((uintptr_t)v_cache % 16 == 0) && ((uintptr_t)qkv % 16 == 0) && cache_stride % 8 == 0,
                "rope_sdpa_decode_batch failed: qkv / K / V must be 16-byte aligned");
    if (scale <= 0.f) scale = 1.0f / sqrtf((float)head_dim);
    RopeSdpaParams p = {};
    p.qkv = (const uint16_t*)qkv; p.q_gamma = (const uint16_t*)q_gamma; p.k_gamma = (const uint16_t*)k_gamma; p.eps = eps;
    p.cos_t = cos_table; p.sin_t = sin_table; p.kc = (uint16_t*)k_cache; p.vc = (uint16_t*)v_cache; p.out = (uint16_t*)out;
    p.hq = n_heads_q; p.n_kv = n_kv; p.max_seq = max_seq; p.pos_buf = pos_buf; p.ctx_buf = ctx_buf;
    p.n_seq = n_seq; p.qkv_stride = (long)(n_heads_q + 2 * n_kv) * 128; p.out_stride = (long)n_heads_q * 128; p.cache_stride = cache_stride;
    p.scale_log2 = scale * 1.4426950408889634f;
    return launch_rope_sdpa(p, (cudaStream_t)stream);
}

size_t pgk_sdpa_workspace_bytes(int, int, int, int) { return kAttnWsBytes; }

}  // extern "C"
