// sampling.cu -- temperature / top-k / top-p token sampling on the device.
//
// Semantics: pygpukit.llm.sampling.sample_token (llm/sampling.py:11-63), the function generate() calls:
//   x = logits / T;  p = softmax(x);  top-k: keep the k largest p, renormalise;  top-p: keep the largest p until their
//   running sum reaches top_p (the crossing element included), renormalise;  T == 0: argmax of what is left, else draw
//   from p in index order (np.random.choice: the first index whose cumulative probability exceeds the uniform number).
// The reference's own CUDA kernels for this (native/ops/sampling/sampling_kernels.cuh:419-760) are not restated: its
// top-k kernel updates a shared top list with unsynchronised atomicExch (the result depends on thread timing) and its
// "top-p" kernel scales the threshold by top_p instead of truncating the distribution.  The uniform number comes from
// the caller (host scalar, or a device buffer for graph replay), as in those kernels' random_val argument.
//
// A cluster of 8 CTAs x 1024 threads (one CTA alone is instruction-bound: 11 us per pass over 152 k logits, 485 us for
// top-k + top-p); the logits (bf16, <= 300 KB) stay in L2 and are re-read by every pass with coalesced 16-byte loads;
// partial results meet over distributed shared memory.  Global warp w (of 256) owns the contiguous index range
// [w * seg, (w + 1) * seg); within it lane l handles elements
// i * 256 + l * 8 .. + 7 of iteration i, so (warp, iteration, lane, element) is index order -- ties and the final
// cumulative walk need that order.  Thresholds are found on the 16-bit order-preserving key of the bf16 logit (the
// probability is a monotone function of it) by a 16-step binary search: counts for top-k, weight sums for top-p.
// Elements tied at a threshold are kept from the highest index down (what a stable ascending argsort, read from its
// end, keeps).  Every reduction has a fixed order: the result is a pure function of the inputs.
#include "common.cuh"
#include "../../include/pgk_b200.h"

#include <climits>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace pgk {

constexpr int kSmpThreads = 1024;
constexpr int kSmpWarps = kSmpThreads / 32;
constexpr int kSmpMaxVocab = 1 << 20;

struct SmpParams {
    const uint16_t* logits;
    int V, seg, iters;          // seg = iters * 256 elements per warp
    float inv_t;                // 1 / temperature (1 when temperature is 0 or 1)
    int greedy;                 // temperature == 0
    int top_k;                  // 0: off
    float top_p;                // >= 1: off
    float u;
    const float* u_buf;         // overrides u when not null
    int32_t* out;
};

__device__ __forceinline__ uint32_t smp_key(uint32_t b) {             // larger logit <=> larger key
    return (b & 0x8000u) ? (~b & 0xffffu) : (b | 0x8000u);
}
__device__ __forceinline__ float smp_key_value(uint32_t k) {          // inverse of smp_key, as f32
    const uint32_t b = (k & 0x8000u) ? (k & 0x7fffu) : (~k & 0xffffu);
    return bf16_bits_to_f32(b);
}

// eight consecutive logits of (warp, iteration, lane); positions >= V read as -inf and are flagged by their index
__device__ __forceinline__ void smp_load8(const SmpParams& p, int base, uint32_t (&b)[8]) {
    if (base + 8 <= p.V) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.logits + base));
        b[0] = v.x & 0xffffu; b[1] = v.x >> 16; b[2] = v.y & 0xffffu; b[3] = v.y >> 16;
        b[4] = v.z & 0xffffu; b[5] = v.z >> 16; b[6] = v.w & 0xffffu; b[7] = v.w >> 16;
    } else {
#pragma unroll
        for (int q = 0; q < 8; ++q) b[q] = (base + q < p.V) ? (uint32_t)__ldg(p.logits + base + q) : 0xff80u;
    }
}

// the two nested filters: element kept iff its key is above the threshold, or equal with index >= cut
struct SmpSet { uint32_t tk1; int cut1; uint32_t tk2; int cut2; };
__device__ __forceinline__ bool smp_in1(const SmpSet& s, uint32_t key, int idx, int V) {
    return idx < V && (key > s.tk1 || (key == s.tk1 && idx >= s.cut1));
}
__device__ __forceinline__ bool smp_in2(const SmpSet& s, uint32_t key, int idx, int V) {
    return smp_in1(s, key, idx, V) && (key > s.tk2 || (key == s.tk2 && idx >= s.cut2));
}

// ---- block reductions with a fixed order (shuffle tree per warp, then a tree over the 32 warp results)
__device__ __forceinline__ float smp_block_sum(float v, float* red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    return warp_sum(red[threadIdx.x & 31]);
}
__device__ __forceinline__ int smp_block_sum_i(int v, float* red) {
    int* r = reinterpret_cast<int*>(red);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) r[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = r[threadIdx.x & 31];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    return t;
}
__device__ __forceinline__ uint32_t smp_block_max_u(uint32_t v, float* red) {
    uint32_t* r = reinterpret_cast<uint32_t*>(red);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) r[threadIdx.x >> 5] = v;
    __syncthreads();
    uint32_t t = r[threadIdx.x & 31];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t = max(t, __shfl_xor_sync(0xffffffffu, t, o));
    return t;
}

// ---- a cluster of kSmpCtas CTAs shares every pass: global warp g = rank * 32 + warp owns [g * seg, (g + 1) * seg).
// A cluster reduction = block reduction, one 32-bit cell per CTA in shared memory, barrier.cluster, every CTA reads
// the kSmpCtas cells over DSMEM in rank order (fixed order: deterministic, and identical in every CTA, so the
// control flow that depends on it stays uniform across the cluster).  The cell alternates between two slots: a CTA
// can only write slot s again after a later cluster barrier, which all CTAs pass after their reads of slot s.
constexpr int kSmpCtas = 8;
constexpr int kSmpGWarps = kSmpCtas * kSmpWarps;

struct SmpCl {
    uint32_t* xch;      // [2] shared
    int par;
};
template <class Op>
__device__ __forceinline__ uint32_t smp_cl_reduce(uint32_t v, SmpCl& cx, Op op) {
    cg::cluster_group cl = cg::this_cluster();
    if (threadIdx.x == 0) cx.xch[cx.par] = v;
    cl.sync();
    uint32_t r = *cl.map_shared_rank(&cx.xch[cx.par], 0);
#pragma unroll
    for (int k = 1; k < kSmpCtas; ++k) r = op(r, *cl.map_shared_rank(&cx.xch[cx.par], k));
    cx.par ^= 1;
    return r;
}
__device__ __forceinline__ float smp_sum(float v, float* red, SmpCl& cx) {
    v = smp_block_sum(v, red);
    return __uint_as_float(smp_cl_reduce(__float_as_uint(v), cx, [](uint32_t a, uint32_t b) {
        return __float_as_uint(__uint_as_float(a) + __uint_as_float(b)); }));
}
__device__ __forceinline__ int smp_sum_i(int v, float* red, SmpCl& cx) {
    v = smp_block_sum_i(v, red);
    return (int)smp_cl_reduce((uint32_t)v, cx, [](uint32_t a, uint32_t b) { return a + b; });
}
__device__ __forceinline__ uint32_t smp_max_u(uint32_t v, float* red, SmpCl& cx) {
    v = smp_block_max_u(v, red);
    return smp_cl_reduce(v, cx, [](uint32_t a, uint32_t b) { return a > b ? a : b; });
}
// value already uniform in the CTA (read from shared memory after a barrier)
__device__ __forceinline__ int smp_cl_max_i(int v, SmpCl& cx) {
    return (int)smp_cl_reduce((uint32_t)v, cx, [](uint32_t a, uint32_t b) { return (int)a > (int)b ? a : b; });
}
__device__ __forceinline__ int smp_cl_min_i(int v, SmpCl& cx) {
    return (int)smp_cl_reduce((uint32_t)v, cx, [](uint32_t a, uint32_t b) { return (int)a < (int)b ? a : b; });
}

__device__ __forceinline__ int smp_gwarp() {
    return (int)cg::this_cluster().block_rank() * kSmpWarps + (int)(threadIdx.x >> 5);
}

// f(key, index) over this thread's elements, in index order.  Four iterations' loads are issued together.
template <class F>
__device__ __forceinline__ void smp_foreach(const SmpParams& p, F f) {
    const int lane = threadIdx.x & 31;
    const int base0 = smp_gwarp() * p.seg + lane * 8;
    for (int i0 = 0; i0 < p.iters; i0 += 4) {
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int base = base0 + (i0 + j) * 256;
            if (i0 + j < p.iters && base + 8 <= p.V) v[j] = __ldg(reinterpret_cast<const uint4*>(p.logits + base));
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int base = base0 + (i0 + j) * 256;
            if (i0 + j >= p.iters || base >= p.V) break;
            uint32_t b[8];
            if (base + 8 <= p.V) {
                b[0] = v[j].x & 0xffffu; b[1] = v[j].x >> 16; b[2] = v[j].y & 0xffffu; b[3] = v[j].y >> 16;
                b[4] = v[j].z & 0xffffu; b[5] = v[j].z >> 16; b[6] = v[j].w & 0xffffu; b[7] = v[j].w >> 16;
            } else {
                smp_load8(p, base, b);
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) f(smp_key(b[q]), base + q);
        }
    }
}

// Index of the `need`-th element, counted from the highest index down, among the candidates with key == c
// (LEVEL 1: every valid element, LEVEL 2: the elements of the first filter).  1 <= need <= number of such elements.
template <int LEVEL>
__device__ int smp_tie_cut(const SmpParams& p, const SmpSet& set, uint32_t c, int need, int* wcnt, int* s_cut, SmpCl& cx) {
    cg::cluster_group cl = cg::this_cluster();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gw = smp_gwarp();
    auto cand = [&](uint32_t key, int idx) {
        return key == c && (LEVEL == 1 ? idx < p.V : smp_in1(set, key, idx, p.V));
    };
    int n = 0;
    smp_foreach(p, [&](uint32_t key, int idx) { n += cand(key, idx) ? 1 : 0; });
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (lane == 0) wcnt[warp] = n;
    if (threadIdx.x == 0) *s_cut = -1;
    cl.sync();
    int above = 0;                                              // candidates in the global warps after this one
    for (int g = gw + 1 + lane; g < kSmpGWarps; g += 32) above += *cl.map_shared_rank(&wcnt[g & 31], g >> 5);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) above += __shfl_xor_sync(0xffffffffu, above, o);
    if (above < need && need <= above + n) {                   // warp-uniform: exactly one warp of the cluster
        int rem = need - above;
        for (int i = p.iters - 1; i >= 0; --i) {               // all 32 lanes stay in the loop: it shuffles
            const int base = gw * p.seg + i * 256 + lane * 8;
            uint32_t b[8];
            smp_load8(p, base, b);
            int c_l = 0;
#pragma unroll
            for (int q = 0; q < 8; ++q) c_l += cand(smp_key(b[q]), base + q) ? 1 : 0;
            int suf = c_l;                                      // candidates in lanes >= this one
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_down_sync(0xffffffffu, suf, o);
                if (lane + o < 32) suf += t;
            }
            int suf_next = __shfl_down_sync(0xffffffffu, suf, 1);
            if (lane == 31) suf_next = 0;
            const int total = __shfl_sync(0xffffffffu, suf, 0);
            if (rem <= total) {
                if (suf_next < rem && rem <= suf) {             // this lane holds it
                    int left = rem - suf_next;
                    for (int q = 7; q >= 0; --q)
                        if (cand(smp_key(b[q]), base + q) && --left == 0) { *s_cut = base + q; break; }
                }
                break;
            }
            rem -= total;
        }
    }
    __syncthreads();
    return smp_cl_max_i(*s_cut, cx);                            // (its barrier also keeps wcnt alive until every CTA has read it)
}

__global__ void __cluster_dims__(kSmpCtas, 1, 1) __launch_bounds__(kSmpThreads, 1) sample_token_kernel(SmpParams p) {
    __shared__ float red[32];
    __shared__ int wcnt[kSmpWarps];
    __shared__ float wtot[kSmpWarps];
    __shared__ uint32_t xch[2];
    __shared__ int s_cut, s_res;
    cg::cluster_group cl = cg::this_cluster();
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gw = smp_gwarp();
    const float u = p.u_buf ? *p.u_buf : p.u;
    SmpCl cx = {xch, 0};
    if (threadIdx.x == 0) s_res = INT_MAX;

    // ---- the largest logit (softmax shift; its weight is exactly 1)
    uint32_t km = 0;
    smp_foreach(p, [&](uint32_t key, int idx) { if (idx < p.V) km = max(km, key); });
    km = smp_max_u(km, red, cx);
    const float lmax = smp_key_value(km);
    const float sc = p.inv_t * 1.4426950408889634f;
    auto weight = [&](uint32_t key) { return exp2f((smp_key_value(key) - lmax) * sc); };

    SmpSet set = {0u, 0, 0u, 0};
    // ---- top-k: the k-th largest key by binary search on counts
    if (p.top_k > 0 && p.top_k < p.V) {
        uint32_t c = 0;
        for (int bit = 15; bit >= 0; --bit) {
            const uint32_t t = c | (1u << bit);
            int n = 0;
            smp_foreach(p, [&](uint32_t key, int idx) { n += (idx < p.V && key >= t) ? 1 : 0; });
            if (smp_sum_i(n, red, cx) >= p.top_k) c = t;
        }
        int ngt = 0;
        smp_foreach(p, [&](uint32_t key, int idx) { ngt += (idx < p.V && key > c) ? 1 : 0; });
        ngt = smp_sum_i(ngt, red, cx);
        const int cut = smp_tie_cut<1>(p, set, c, p.top_k - ngt, wcnt, &s_cut, cx);
        set.tk1 = c; set.cut1 = cut;
    }
    // ---- top-p: the largest key whose tail sum reaches top_p of the kept mass
    if (p.top_p < 1.f) {
        float z1 = 0.f;
        smp_foreach(p, [&](uint32_t key, int idx) { if (smp_in1(set, key, idx, p.V)) z1 += weight(key); });
        z1 = smp_sum(z1, red, cx);
        const float target = fmaxf(p.top_p * z1, 1e-30f);
        uint32_t c = 0;
        for (int bit = 15; bit >= 0; --bit) {
            const uint32_t t = c | (1u << bit);
            float f = 0.f;
            smp_foreach(p, [&](uint32_t key, int idx) { if (key >= t && smp_in1(set, key, idx, p.V)) f += weight(key); });
            if (smp_sum(f, red, cx) >= target) c = t;
        }
        float sgt = 0.f;
        int ties = 0;
        smp_foreach(p, [&](uint32_t key, int idx) {
            if (smp_in1(set, key, idx, p.V)) { if (key > c) sgt += weight(key); else if (key == c) ++ties; }
        });
        sgt = smp_sum(sgt, red, cx);
        ties = smp_sum_i(ties, red, cx);
        int need = (int)ceilf((target - sgt) / weight(c));
        need = min(max(need, 1), max(ties, 1));
        const int cut = smp_tie_cut<2>(p, set, c, need, wcnt, &s_cut, cx);
        set.tk2 = c; set.cut2 = cut;
    }
    // ---- the token
    if (p.greedy) {
        int best = INT_MAX;
        smp_foreach(p, [&](uint32_t key, int idx) { if (key == km && smp_in2(set, key, idx, p.V)) best = min(best, idx); });
        if (best != INT_MAX) atomicMin(&s_res, best);
        __syncthreads();
        const int res = smp_cl_min_i(s_res, cx);
        if (cl.block_rank() == 0 && threadIdx.x == 0) p.out[0] = res == INT_MAX ? 0 : res;
        cl.sync();                                              // no CTA leaves while its cells may still be read
        return;
    }
    // cumulative walk in index order: per-warp totals, a fixed-order prefix over the 256 global warps, then every warp
    // walks its own range and reports the first index whose running sum exceeds u * total (the smallest one wins)
    float z = 0.f;
    smp_foreach(p, [&](uint32_t key, int idx) { if (smp_in2(set, key, idx, p.V)) z += weight(key); });
    z = warp_sum(z);
    if (lane == 0) wtot[warp] = z;
    cl.sync();
    float before = 0.f, total = 0.f;
    for (int g = lane; g < kSmpGWarps; g += 32) {
        const float t = *cl.map_shared_rank(&wtot[g & 31], g >> 5);
        total += t;
        if (g < gw) before += t;
    }
    total = warp_sum(total);
    float run0 = warp_sum(before);
    const float tau = u * total;
    for (int i = 0; i < p.iters; ++i) {                         // all 32 lanes stay in the loop: it shuffles
        const int base = gw * p.seg + i * 256 + lane * 8;
        uint32_t b[8];
        smp_load8(p, base, b);
        float w8[8], s_l = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const uint32_t key = smp_key(b[q]);
            w8[q] = smp_in2(set, key, base + q, p.V) ? weight(key) : 0.f;
            s_l += w8[q];
        }
        float pre = s_l;                                        // inclusive prefix over the lanes
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float t = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += t;
        }
        bool found = false;
        if (run0 + pre > tau) {                                 // the crossing is in this lane or an earlier one
            float r = run0 + (pre - s_l);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                r += w8[q];
                if (!found && w8[q] > 0.f && r > tau) { atomicMin(&s_res, base + q); found = true; }
            }
        }
        run0 += __shfl_sync(0xffffffffu, pre, 31);
        if (__any_sync(0xffffffffu, found)) break;
    }
    __syncthreads();
    int res = smp_cl_min_i(s_res, cx);
    if (res == INT_MAX) {                                       // u * total rounded up to the total: the last kept index
        int last = -1;
        smp_foreach(p, [&](uint32_t key, int idx) { if (smp_in2(set, key, idx, p.V) && weight(key) > 0.f) last = max(last, idx); });
        if (threadIdx.x == 0) s_cut = -1;                       // (s_cut is free here: reused as the max cell)
        __syncthreads();
        if (last >= 0) atomicMax(&s_cut, last);
        __syncthreads();
        res = max(smp_cl_max_i(s_cut, cx), 0);
    }
    if (cl.block_rank() == 0 && threadIdx.x == 0) p.out[0] = res;
    cl.sync();                                                  // no CTA leaves while its cells may still be read
}

}  // namespace pgk

using namespace pgk;

extern "C" {

/* sample_token (llm/sampling.py:11-63) on the device: logits bf16 [vocab_size] -> token_out[0].  random_buf (device,
 * one float in [0, 1)) overrides random_val when not NULL (graph replay: update the buffer, replay). */
int pgk_sample_token_bf16(const void* logits, int vocab_size, float temperature, int top_k, float top_p,
                          float random_val, const float* random_buf, int32_t* token_out, void* stream) {
    PGK_REQUIRE(logits && token_out && vocab_size >= 1 && vocab_size <= kSmpMaxVocab,
                "sample_token failed: needs logits, an output buffer and 1 <= vocab_size <= %d (got %d)", kSmpMaxVocab, vocab_size);
    PGK_REQUIRE(temperature >= 0.f, "sample_token failed: temperature must be >= 0");
    PGK_REQUIRE(((uintptr_t)logits % 16) == 0, "sample_token failed: logits must be 16-byte aligned");
    PGK_REQUIRE(random_buf || (random_val >= 0.f && random_val < 1.f), "sample_token failed: random_val must be in [0, 1)");
    SmpParams p = {};
    p.logits = (const uint16_t*)logits; p.V = vocab_size;
    p.iters = (vocab_size + kSmpGWarps * 256 - 1) / (kSmpGWarps * 256);
    p.seg = p.iters * 256;
    p.greedy = temperature == 0.f;
    p.inv_t = (temperature != 1.f && temperature > 0.f) ? 1.f / temperature : 1.f;
    p.top_k = top_k; p.top_p = top_p; p.u = random_val; p.u_buf = random_buf; p.out = token_out;
    PGK_CUDA(launch(sample_token_kernel, dim3(kSmpCtas), dim3(kSmpThreads), 0, (cudaStream_t)stream, p));
    return 0;
}

}  // extern "C"
