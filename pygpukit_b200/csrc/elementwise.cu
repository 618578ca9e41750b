// elementwise.cu -- the small ops that bracket the GEMVs in a decode step: RMSNorm /
// LayerNorm, RoPE, SiLU/GELU, add/mul, bias, cast, embedding gather, argmax, transpose,
// KV-cache write, and the fused QK-norm + RoPE + KV-write bracket.
// All f32 math with the reference's rounding points; file:line
// citations of the reference kernels are on each op.
#include "common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

template <int DT> struct Elem;
template <> struct Elem<0> {
    using T = float;
    static __device__ __forceinline__ float ld(const void* p, size_t i) { return ((const float*)p)[i]; }
    static __device__ __forceinline__ void st(void* p, size_t i, float v) { ((float*)p)[i] = v; }
};
template <> struct Elem<1> {
    using T = __half;
    static __device__ __forceinline__ float ld(const void* p, size_t i) { return __half2float(((const __half*)p)[i]); }
    static __device__ __forceinline__ void st(void* p, size_t i, float v) { ((__half*)p)[i] = __float2half_rn(v); }
};
template <> struct Elem<2> {
    using T = __nv_bfloat16;
    static __device__ __forceinline__ float ld(const void* p, size_t i) { return __bfloat162float(((const __nv_bfloat16*)p)[i]); }
    static __device__ __forceinline__ void st(void* p, size_t i, float v) { ((__nv_bfloat16*)p)[i] = __float2bfloat16_rn(v); }
};

#define PGK_DISPATCH_DT(dt, NAME, ...)                                           \
    switch (dt) {                                                                \
        case 0: { constexpr int DT = 0; __VA_ARGS__; break; }                    \
        case 1: { constexpr int DT = 1; __VA_ARGS__; break; }                    \
        case 2: { constexpr int DT = 2; __VA_ARGS__; break; }                    \
        default: pgk::set_error(NAME " failed: unsupported dtype code %d", dt); return 1; \
    }

__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    // the reference folds the per-warp totals with a second shfl_down tree in warp 0
    // (norm_kernels.cuh:553-570); a butterfly over {red[lane] | 0} gives lane 0's pairing everywhere
    return warp_sum(lane < nw ? red[lane] : 0.f);
}

// ---------------------------------------------------------------- RMSNorm / LayerNorm
// rmsnorm_*_kernel norm_kernels.cuh:375-585: y = x * rsqrt(mean(x^2)+eps) * gamma (left to right).
template <int DT, bool RESIDUAL>
__global__ void rmsnorm_kernel(const void* __restrict__ x, const void* __restrict__ res,
                               const void* __restrict__ gamma, void* __restrict__ y, int features, float eps) {
    __shared__ float red[32];
    griddep_wait();
    griddep_launch();
    const size_t base = (size_t)blockIdx.x * features;
    // thread t owns elements t, t + bs, ... (the reference's partial sums); up to 32 of them are loaded
    // up front (all loads in flight at once) and kept in registers for the second pass
    constexpr int MAXE = 32;
    float xv[MAXE];
    const int bs = blockDim.x;
    const bool cached = features <= bs * MAXE;
    float ss = 0.f;
    if (cached) {
#pragma unroll
        for (int e = 0; e < MAXE; ++e) {
            const int i = threadIdx.x + e * bs;
            xv[e] = 0.f;
            if (i < features) {
                xv[e] = Elem<DT>::ld(x, base + i);
                if (RESIDUAL) xv[e] += Elem<DT>::ld(res, base + i);
            }
        }
#pragma unroll
        for (int e = 0; e < MAXE; ++e)
            if (threadIdx.x + e * bs < features) ss = fmaf(xv[e], xv[e], ss);
    } else {
        for (int i = threadIdx.x; i < features; i += bs) {
            float v = Elem<DT>::ld(x, base + i);
            if (RESIDUAL) v += Elem<DT>::ld(res, base + i);   // fused_kernels.cuh:36-220 keeps the sum in f32
            ss = fmaf(v, v, ss);
        }
    }
    float tot = block_sum(ss, red);
    float inv = rsqrtf(tot / (float)features + eps);
    if (cached) {
#pragma unroll
        for (int e = 0; e < MAXE; ++e) {
            const int i = threadIdx.x + e * bs;
            if (i < features) Elem<DT>::st(y, base + i, xv[e] * inv * Elem<DT>::ld(gamma, i));
        }
    } else {
        for (int i = threadIdx.x; i < features; i += bs) {
            float v = Elem<DT>::ld(x, base + i);
            if (RESIDUAL) v += Elem<DT>::ld(res, base + i);
            Elem<DT>::st(y, base + i, v * inv * Elem<DT>::ld(gamma, i));
        }
    }
}

// layernorm_*_kernel norm_kernels.cuh:32-330: mean, var = mean((x-mean)^2), (x-mean)*rsqrt(var+eps)*g+b
template <int DT>
__global__ void layernorm_kernel(const void* __restrict__ x, const void* __restrict__ gamma,
                                 const void* __restrict__ beta, void* __restrict__ y, int features, float eps) {
    __shared__ float red[32];
    griddep_wait();
    griddep_launch();
    const size_t base = (size_t)blockIdx.x * features;
    float s = 0.f;
    for (int i = threadIdx.x; i < features; i += blockDim.x) s += Elem<DT>::ld(x, base + i);
    float mean = block_sum(s, red) / (float)features;
    float vs = 0.f;
    for (int i = threadIdx.x; i < features; i += blockDim.x) {
        float d = Elem<DT>::ld(x, base + i) - mean;
        vs = fmaf(d, d, vs);
    }
    float var = block_sum(vs, red) / (float)features;
    float inv = rsqrtf(var + eps);
    for (int i = threadIdx.x; i < features; i += blockDim.x) {
        float v = (Elem<DT>::ld(x, base + i) - mean) * inv * Elem<DT>::ld(gamma, i);
        if (beta) v += Elem<DT>::ld(beta, i);
        Elem<DT>::st(y, base + i, v);
    }
}

// ---------------------------------------------------------------- RoPE (NeoX half-split)
// rope_*_kernel / rope_*_f32table_kernel elementwise_kernels.cuh:82-366.  One thread per
// (position, head, d < hd/2) of q then k.
template <int DT, int TDT>
__global__ void rope_kernel(void* __restrict__ q, void* __restrict__ k, const void* __restrict__ cs,
                            const void* __restrict__ sn, int seq, int hq, int hk, int hd) {
    griddep_wait();
    griddep_launch();
    const int half = hd >> 1;
    const long nq = (long)seq * hq * half, nk = (long)seq * hk * half;
    long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq + nk) return;
    void* buf = q;
    int heads = hq;
    if (t >= nq) { t -= nq; buf = k; heads = hk; }
    const int d = (int)(t % half);
    const long sh = t / half;          // s * heads + h
    const int s = (int)(sh / heads);
    const size_t o = (size_t)sh * hd + d;
    float c = Elem<TDT>::ld(cs, (size_t)s * hd + d), sv = Elem<TDT>::ld(sn, (size_t)s * hd + d);
    float x0 = Elem<DT>::ld(buf, o), x1 = Elem<DT>::ld(buf, o + half);
    Elem<DT>::st(buf, o, x0 * c - x1 * sv);
    Elem<DT>::st(buf, o + half, x1 * c + x0 * sv);
}

// ---------------------------------------------------------------- unary / binary glue
enum { OP_SILU = 0, OP_GELU = 1 };
enum { OP_ADD = 0, OP_MUL = 1, OP_SWIGLU = 2, OP_SILU_MUL = 3 };

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }   // activation_kernels.cuh:40-42
__device__ __forceinline__ float gelu_f(float x) {                                       // activation_kernels.cuh:26-31
    return x * 0.5f * (1.0f + tanhf(0.7978845608f * (x + 0.044715f * x * x * x)));
}

template <int DT, int OP>
__global__ void unary_kernel(const void* __restrict__ x, void* __restrict__ y, size_t n) {
    griddep_wait();
    griddep_launch();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = Elem<DT>::ld(x, i);
    Elem<DT>::st(y, i, OP == OP_SILU ? silu_f(v) : gelu_f(v));
}

// add: f32 add then one rounding (elementwise/elementwise_kernels.cuh:55-60);
// mul: exact f32 product then rounding; swiglu: silu_f32(g)*u one rounding (fused_kernels.cuh:233-247)
template <int DT, int OP>
__global__ void binary_kernel(const void* __restrict__ a, const void* __restrict__ b, void* __restrict__ c, size_t n) {
    griddep_wait();
    griddep_launch();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float x = Elem<DT>::ld(a, i), y = Elem<DT>::ld(b, i);
    if (OP == OP_SILU_MUL) {          // silu -> store rounding -> native multiply: the two-launch form's two roundings
        Elem<DT>::st(c, i, silu_f(x));
        Elem<DT>::st(c, i, Elem<DT>::ld(c, i) * y);
        return;
    }
    Elem<DT>::st(c, i, OP == OP_ADD ? x + y : (OP == OP_MUL ? x * y : silu_f(x) * y));
}

// in-place native-precision add/mul (__hadd/__hmul: one rounding of the exact result),
// elementwise_kernels.cuh:370-470
template <int DT, int OP>
__global__ void binary_inplace_kernel(void* __restrict__ a, const void* __restrict__ b, size_t n) {
    griddep_wait();
    griddep_launch();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    using T = typename Elem<DT>::T;
    T* pa = (T*)a;
    const T* pb = (const T*)b;
    if constexpr (DT == 0) pa[i] = (OP == OP_ADD) ? pa[i] + pb[i] : pa[i] * pb[i];
    else pa[i] = (OP == OP_ADD) ? __hadd(pa[i], pb[i]) : __hmul(pa[i], pb[i]);
}

// bias_add_*_kernel elementwise_kernels.cuh:22-75: f32 add, one rounding
template <int DT>
__global__ void bias_add_kernel(void* __restrict__ x, const void* __restrict__ bias, size_t n, int features) {
    griddep_wait();
    griddep_launch();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Elem<DT>::st(x, i, Elem<DT>::ld(x, i) + Elem<DT>::ld(bias, i % features));
}

template <int SDT, int DDT>
__global__ void cast_kernel(const void* __restrict__ s, void* __restrict__ d, size_t n) {
    griddep_wait();
    griddep_launch();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) Elem<DDT>::st(d, i, Elem<SDT>::ld(s, i));
}

// GPT-2 input: hidden[t] = wte[token[t]] + wpe[pos + t], one rounding (models/causal.py:137-146 adds the
// two rows on the host in the storage dtype); token ids and the position may live in device buffers.
template <int DT>
__global__ void embedding_pos_kernel(const void* __restrict__ wte, const void* __restrict__ wpe, void* __restrict__ out,
                                     int H, const int32_t* __restrict__ token_buf, int n_tokens, int pos,
                                     const int32_t* __restrict__ pos_buf) {
    griddep_wait();
    griddep_launch();
    const int p0 = pos_buf ? *pos_buf : pos;
    for (int t = blockIdx.y; t < n_tokens; t += gridDim.y) {
        const size_t tok = (size_t)token_buf[t] * H, pr = (size_t)(p0 + t) * H, o = (size_t)t * H;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H; i += gridDim.x * blockDim.x)
            Elem<DT>::st(out, o + i, Elem<DT>::ld(wte, tok + i) + Elem<DT>::ld(wpe, pr + i));
    }
}

// ---------------------------------------------------------------- gather / argmax / layout
// embedding_lookup(_ptr,_batch) embedding_kernels.cuh:27-165: row copy; token id from host or device
__global__ void embedding_kernel(const uint8_t* __restrict__ table, uint8_t* __restrict__ out, size_t row_bytes,
                                 int token, const int32_t* __restrict__ token_buf, int n_tokens) {
    griddep_wait();
    griddep_launch();
    const size_t vec = row_bytes >> 4;
    for (int t = blockIdx.y; t < n_tokens; t += gridDim.y) {
        long tok = token_buf ? (long)token_buf[t] : (long)token;
        const uint4* src = reinterpret_cast<const uint4*>(table + (size_t)tok * row_bytes);
        uint4* dst = reinterpret_cast<uint4*>(out + (size_t)t * row_bytes);
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < vec; i += (size_t)gridDim.x * blockDim.x)
            dst[i] = __ldg(src + i);
        if (blockIdx.x == 0)
            for (size_t i = (vec << 4) + threadIdx.x; i < row_bytes; i += blockDim.x)
                out[(size_t)t * row_bytes + i] = table[(size_t)tok * row_bytes + i];
    }
}

// sample_argmax_*_kernel sampling_kernels.cuh:55-200: first maximum wins (np.argmax)
template <int DT>
__global__ void argmax_kernel(const void* __restrict__ x, int32_t* __restrict__ out, int n) {
    __shared__ float sv[32];
    __shared__ int si[32];
    griddep_wait();
    griddep_launch();
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float v = Elem<DT>::ld(x, i);
        if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)((blockDim.x + 31) >> 5); ++w)
            if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
        *out = (bi == 0x7fffffff) ? 0 : bi;
    }
}
// gate/up columns interleaved in groups of 8 ([g0..g7, u0..u7, g8..], the fused NVF4 MLP tensor): out[m, j] =
// bf16( bf16(silu(g)) * u ) -- silu_bf16_kernel then mul_inplace_bf16_kernel of the reference (activation_kernels.cuh:164-171,
// elementwise_kernels.cuh:429-438), two roundings
__global__ void swiglu_interleaved8_kernel(const uint16_t* __restrict__ gu, uint16_t* __restrict__ out, int M, int I) {
    griddep_wait();
    griddep_launch();
    const size_t total = (size_t)M * I;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t m = i / I;
        const int j = (int)(i - m * I);
        const uint16_t* row = gu + m * (size_t)(2 * I) + (size_t)(j >> 3) * 16 + (j & 7);
        const float g = bf16_bits_to_f32(row[0]), u = bf16_bits_to_f32(row[8]);
        const float sl = round_bf16(g / (1.0f + __expf(-g)));
        out[i] = f32_to_bf16_bits(sl * u);
    }
}
// one CTA per row of a [rows, n] matrix (greedy tokens of a decode batch); out2 (optional) receives the same tokens
template <int DT>
__global__ void argmax_rows_kernel(const void* __restrict__ x, int32_t* __restrict__ out, int32_t* __restrict__ out2, int n) {
    __shared__ float sv[32];
    __shared__ int si[32];
    griddep_wait();
    griddep_launch();
    const size_t base = (size_t)blockIdx.x * n;
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float v = Elem<DT>::ld(x, base + i);
        if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)((blockDim.x + 31) >> 5); ++w)
            if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
        const int t = (bi == 0x7fffffff) ? 0 : bi;
        out[blockIdx.x] = t;
        if (out2) out2[blockIdx.x] = t;
    }
}

// transpose_3d_021: [d0,d1,d2] -> [d1,d0,d2]
template <typename T>
__global__ void transpose021_kernel(const T* __restrict__ in, T* __restrict__ out, int d0, int d1, int d2) {
    griddep_wait();
    griddep_launch();
    size_t n = (size_t)d0 * d1 * d2;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = (int)(i % d2);
    size_t r = i / d2;
    int b = (int)(r % d1), a = (int)(r / d1);
    out[((size_t)b * d0 + a) * d2 + c] = in[i];
}

// kv_cache_update(_gqa)(_ptr) / kv_cache_prefill_gqa kv_cache_kernels.cuh:176-421.
// new_kv [seq, n_kv, hd] -> cache [heads_out, max_seq, hd]; heads_out = G * n_kv replicates.
__global__ void kv_update_kernel(const uint8_t* __restrict__ new_kv, uint8_t* __restrict__ cache, int seq, int n_kv,
                                 int heads_out, size_t row_bytes, int max_seq, int pos,
                                 const int32_t* __restrict__ pos_buf) {
    griddep_wait();
    griddep_launch();
    const int p0 = pos_buf ? *pos_buf : pos;
    const int G = heads_out / n_kv;
    const size_t vec = row_bytes >> 4;                       // head_dim * elem is a multiple of 16 in practice
    const size_t total = (size_t)seq * heads_out * vec;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t v = i % vec;
        size_t r = i / vec;
        int h = (int)(r % heads_out), s = (int)(r / heads_out);
        if (p0 + s >= max_seq) continue;
        const uint4* src = reinterpret_cast<const uint4*>(new_kv + ((size_t)s * n_kv + h / G) * row_bytes);
        uint4* dst = reinterpret_cast<uint4*>(cache + ((size_t)h * max_seq + p0 + s) * row_bytes);
        dst[v] = src[v];
    }
}
__global__ void kv_update_bytes_kernel(const uint8_t* __restrict__ new_kv, uint8_t* __restrict__ cache, int seq,
                                       int n_kv, int heads_out, size_t row_bytes, int max_seq, int pos,
                                       const int32_t* __restrict__ pos_buf) {
    griddep_wait();
    griddep_launch();
    const int p0 = pos_buf ? *pos_buf : pos;
    const int G = heads_out / n_kv;
    const size_t total = (size_t)seq * heads_out * row_bytes;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t v = i % row_bytes;
        size_t r = i / row_bytes;
        int h = (int)(r % heads_out), s = (int)(r / heads_out);
        if (p0 + s >= max_seq) continue;
        cache[((size_t)h * max_seq + p0 + s) * row_bytes + v] = new_kv[((size_t)s * n_kv + h / G) * row_bytes + v];
    }
}

// Fused decode bracket: per (row m, head) one warp does [QK-RMSNorm] -> RoPE(f32 table, position
// pos+m) -> in-place store (q, k) and K/V cache write.  Rounds to bf16 after the norm and after
// the rotation exactly like rmsnorm -> rope_inplace_f32table -> kv_cache_update_gqa
// (layers/attention.py:326-351).
__global__ void qk_norm_rope_kv_kernel(uint16_t* __restrict__ qkv, const uint16_t* __restrict__ qg,
                                       const uint16_t* __restrict__ kg, float eps, const float* __restrict__ cs,
                                       const float* __restrict__ sn, uint16_t* __restrict__ kc,
                                       uint16_t* __restrict__ vc, int M, int hq, int hkv, int hd, int max_seq,
                                       int pos, const int32_t* __restrict__ pos_buf) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31;
    const int gw = (int)(((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int H = hq + 2 * hkv;
    if (gw >= M * H) return;
    const int m = gw / H, h = gw % H;
    const int p = (pos_buf ? *pos_buf : pos) + m;
    uint16_t* row = qkv + ((size_t)m * H + h) * hd;
    const int half = hd >> 1;
    if (h >= hq + hkv) {   // V head: straight copy into the cache
        if (p < max_seq) {
            uint16_t* dst = vc + ((size_t)(h - hq - hkv) * max_seq + p) * hd;
            for (int d = lane; d < hd; d += 32) dst[d] = row[d];
        }
        return;
    }
    const bool is_k = h >= hq;
    const uint16_t* gamma = is_k ? kg : qg;
    float inv = 1.f;
    if (gamma) {
        // same summation tree as the reference's rmsnorm block of bs = min(256, hd^32) threads
        // (virtual thread t = lane + 32 w owns elements t, t + bs, ...): per-warp trees, then a tree
        // over the warp totals -- so the fused bracket rounds exactly like the separate QK-norm launch
        const int bs = min(256, (hd + 31) & ~31), nw = bs >> 5;
        float tot = 0.f;
        for (int w = 0; w < nw; ++w) {
            float part = 0.f;
            for (int d = lane + 32 * w; d < hd; d += bs) { float v = bf16_bits_to_f32(row[d]); part = fmaf(v, v, part); }
            part = warp_sum(part);
            if (lane == w) tot = part;
        }
        float ss = warp_sum(tot);
        inv = rsqrtf(ss / (float)hd + eps);
    }
    uint16_t* dstk = (is_k && p < max_seq) ? kc + ((size_t)(h - hq) * max_seq + p) * hd : nullptr;
    for (int d = lane; d < half; d += 32) {
        float x0 = bf16_bits_to_f32(row[d]), x1 = bf16_bits_to_f32(row[d + half]);
        if (gamma) {
            x0 = round_bf16(x0 * inv * bf16_bits_to_f32(gamma[d]));
            x1 = round_bf16(x1 * inv * bf16_bits_to_f32(gamma[d + half]));
        }
        uint16_t y0, y1;
        if (cs) {
            float c = cs[(size_t)p * hd + d], s = sn[(size_t)p * hd + d];
            y0 = f32_to_bf16_bits(x0 * c - x1 * s); y1 = f32_to_bf16_bits(x1 * c + x0 * s);
        } else {                                  // no rotary embedding (GPT-2): plain K/V cache write
            y0 = f32_to_bf16_bits(x0); y1 = f32_to_bf16_bits(x1);
        }
        row[d] = y0; row[d + half] = y1;
        if (dstk) { dstk[d] = y0; dstk[d + half] = y1; }
    }
}

static inline unsigned blocks_for(size_t n, int bs) { return (unsigned)((n + bs - 1) / bs); }

}  // namespace pgk

using namespace pgk;
#define ST(s) ((cudaStream_t)(s))

extern "C" {

int pgk_rmsnorm(const void* x, const void* gamma, void* y, int rows, int features, float eps, int dtype, void* stream) {
    PGK_REQUIRE(rows >= 0 && features > 0, "rmsnorm failed: invalid shape");
    if (rows == 0) return 0;
    int bs = (features + 31) / 32 * 32;          // the reference's block size (rmsnorm.inl:18-19): same partial sums
    if (bs > 256) bs = 256;
    PGK_DISPATCH_DT(dtype, "rmsnorm",
        PGK_CUDA(launch(rmsnorm_kernel<DT, false>, dim3(rows), dim3(bs), 0, ST(stream), x, (const void*)nullptr, gamma, y, features, eps)));
    return 0;
}
int pgk_rmsnorm_residual(const void* x, const void* residual, const void* gamma, void* y, int rows, int features,
                         float eps, int dtype, void* stream) {
    PGK_REQUIRE(rows >= 0 && features > 0, "rmsnorm_residual failed: invalid shape");
    if (rows == 0) return 0;
    PGK_DISPATCH_DT(dtype, "rmsnorm_residual",
        PGK_CUDA(launch(rmsnorm_kernel<DT, true>, dim3(rows), dim3(256), 0, ST(stream), x, residual, gamma, y, features, eps)));
    return 0;
}
int pgk_layernorm(const void* x, const void* gamma, const void* beta, void* y, int rows, int features, float eps,
                  int dtype, void* stream) {
    PGK_REQUIRE(rows >= 0 && features > 0, "layernorm failed: invalid shape");
    if (rows == 0) return 0;
    PGK_DISPATCH_DT(dtype, "layernorm",
        PGK_CUDA(launch(layernorm_kernel<DT>, dim3(rows), dim3(256), 0, ST(stream), x, gamma, beta, y, features, eps)));
    return 0;
}
int pgk_rope_inplace(void* q, void* k, const void* cos, const void* sin, int seq, int n_heads_q, int n_heads_k,
                     int head_dim, int dtype, int table_dtype, void* stream) {
    PGK_REQUIRE(head_dim % 2 == 0, "rope_inplace failed: head_dim must be even");
    size_t n = (size_t)seq * (n_heads_q + n_heads_k) * (head_dim / 2);
    if (n == 0) return 0;
    const bool f32tab = (table_dtype == 0);
    PGK_DISPATCH_DT(dtype, "rope_inplace",
        if (f32tab) PGK_CUDA(launch(rope_kernel<DT, 0>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), q, k, cos, sin, seq, n_heads_q, n_heads_k, head_dim));
        else        PGK_CUDA(launch(rope_kernel<DT, DT>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), q, k, cos, sin, seq, n_heads_q, n_heads_k, head_dim)));
    return 0;
}
int pgk_silu(const void* x, void* y, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "silu", PGK_CUDA(launch(unary_kernel<DT, OP_SILU>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), x, y, n)));
    return 0;
}
int pgk_gelu(const void* x, void* y, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "gelu", PGK_CUDA(launch(unary_kernel<DT, OP_GELU>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), x, y, n)));
    return 0;
}
/* y = round(round(silu(gate)) * up): silu + mul_inplace of the layer path (two roundings) in one launch */
int pgk_silu_mul(const void* gate, const void* up, void* y, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "silu_mul", PGK_CUDA(launch(binary_kernel<DT, OP_SILU_MUL>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), gate, up, y, n)));
    return 0;
}
int pgk_swiglu(const void* gate, const void* up, void* y, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "swiglu", PGK_CUDA(launch(binary_kernel<DT, OP_SWIGLU>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), gate, up, y, n)));
    return 0;
}
int pgk_add(const void* a, const void* b, void* c, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "add", PGK_CUDA(launch(binary_kernel<DT, OP_ADD>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), a, b, c, n)));
    return 0;
}
int pgk_mul(const void* a, const void* b, void* c, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "mul", PGK_CUDA(launch(binary_kernel<DT, OP_MUL>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), a, b, c, n)));
    return 0;
}
int pgk_add_inplace(void* a, const void* b, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "add_inplace", PGK_CUDA(launch(binary_inplace_kernel<DT, OP_ADD>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), a, b, n)));
    return 0;
}
int pgk_mul_inplace(void* a, const void* b, size_t n, int dtype, void* stream) {
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "mul_inplace", PGK_CUDA(launch(binary_inplace_kernel<DT, OP_MUL>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), a, b, n)));
    return 0;
}
int pgk_bias_add_inplace(void* x, const void* bias, int rows, int features, int dtype, void* stream) {
    size_t n = (size_t)rows * features;
    if (n == 0) return 0;
    PGK_DISPATCH_DT(dtype, "bias_add_inplace", PGK_CUDA(launch(bias_add_kernel<DT>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), x, bias, n, features)));
    return 0;
}
int pgk_cast(const void* src, void* dst, size_t n, int src_dtype, int dst_dtype, void* stream) {
    if (n == 0) return 0;
    unsigned g = blocks_for(n, 256);
#define PGK_CAST(S, D) if (src_dtype == S && dst_dtype == D) { PGK_CUDA(launch(cast_kernel<S, D>, dim3(g), dim3(256), 0, ST(stream), src, dst, n)); return 0; }
    PGK_CAST(0, 1) PGK_CAST(0, 2) PGK_CAST(1, 0) PGK_CAST(2, 0) PGK_CAST(1, 2) PGK_CAST(2, 1)
    PGK_CAST(0, 0) PGK_CAST(1, 1) PGK_CAST(2, 2)
#undef PGK_CAST
    set_error("cast failed: unsupported dtype pair %d -> %d", src_dtype, dst_dtype);
    return 1;
}
int pgk_embedding_lookup(const void* table, void* out, int hidden, int token, int elem_bytes, void* stream) {
    size_t rb = (size_t)hidden * elem_bytes;
    PGK_CUDA(launch(embedding_kernel, dim3(blocks_for(rb / 16 + 1, 256), 1), dim3(256), 0, ST(stream),
                    (const uint8_t*)table, (uint8_t*)out, rb, token, (const int32_t*)nullptr, 1));
    return 0;
}
int pgk_embedding_lookup_ptr(const void* table, void* out, int hidden, const int32_t* token_buf, int n_tokens,
                             int elem_bytes, void* stream) {
    if (n_tokens <= 0) return 0;
    size_t rb = (size_t)hidden * elem_bytes;
    PGK_CUDA(launch(embedding_kernel, dim3(blocks_for(rb / 16 + 1, 256), n_tokens > 65535 ? 65535 : n_tokens), dim3(256), 0,
                    ST(stream), (const uint8_t*)table, (uint8_t*)out, rb, 0, token_buf, n_tokens));
    return 0;
}
int pgk_embedding_lookup_pos(const void* wte, const void* wpe, void* out, int hidden, const int32_t* token_buf,
                             int n_tokens, int pos, const int32_t* pos_buf, int dtype, void* stream) {
    if (n_tokens <= 0) return 0;
    PGK_REQUIRE(token_buf != nullptr, "embedding_lookup_pos failed: token_buf is NULL");
    PGK_DISPATCH_DT(dtype, "embedding_lookup_pos",
                    PGK_CUDA(launch(embedding_pos_kernel<DT>, dim3(blocks_for(hidden, 256), n_tokens > 65535 ? 65535 : n_tokens),
                                    dim3(256), 0, ST(stream), wte, wpe, out, hidden, token_buf, n_tokens, pos, pos_buf)));
    return 0;
}
int pgk_argmax(const void* logits, int32_t* out, int n, int dtype, void* stream) {
    PGK_REQUIRE(n > 0, "argmax failed: empty input");
    PGK_DISPATCH_DT(dtype, "argmax", PGK_CUDA(launch(argmax_kernel<DT>, dim3(1), dim3(1024), 0, ST(stream), logits, out, n)));
    return 0;
}
int pgk_swiglu_interleaved8_bf16(const void* gate_up, void* out, int M, int I, void* stream) {
    PGK_REQUIRE(M >= 1 && I >= 8 && I % 8 == 0, "swiglu_interleaved8 failed: invalid shape M=%d I=%d", M, I);
    PGK_CUDA(launch(swiglu_interleaved8_kernel, dim3(blocks_for((size_t)M * I, 256)), dim3(256), 0, ST(stream),
                    (const uint16_t*)gate_up, (uint16_t*)out, M, I));
    return 0;
}
int pgk_argmax_rows(const void* logits, int32_t* out, int32_t* out2, int rows, int n, int dtype, void* stream) {
    PGK_REQUIRE(n > 0 && rows >= 1, "argmax_rows failed: empty input");
    PGK_DISPATCH_DT(dtype, "argmax_rows", PGK_CUDA(launch(argmax_rows_kernel<DT>, dim3(rows), dim3(1024), 0, ST(stream), logits, out, out2, n)));
    return 0;
}
int pgk_transpose_021(const void* in, void* out, int d0, int d1, int d2, int elem_bytes, void* stream) {
    size_t n = (size_t)d0 * d1 * d2;
    if (n == 0) return 0;
    if (elem_bytes == 2) PGK_CUDA(launch(transpose021_kernel<uint16_t>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), (const uint16_t*)in, (uint16_t*)out, d0, d1, d2));
    else if (elem_bytes == 4) PGK_CUDA(launch(transpose021_kernel<uint32_t>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), (const uint32_t*)in, (uint32_t*)out, d0, d1, d2));
    else if (elem_bytes == 1) PGK_CUDA(launch(transpose021_kernel<uint8_t>, dim3(blocks_for(n, 256)), dim3(256), 0, ST(stream), (const uint8_t*)in, (uint8_t*)out, d0, d1, d2));
    else { set_error("transpose_3d_021 failed: unsupported element size %d", elem_bytes); return 1; }
    return 0;
}
int pgk_kv_cache_update(const void* new_kv, void* cache, int seq, int n_kv, int num_heads_out, int head_dim,
                        int max_seq, int pos, const int32_t* pos_buf, int elem_bytes, void* stream) {
    PGK_REQUIRE(n_kv > 0 && num_heads_out % n_kv == 0, "kv_cache_update failed: num_heads %d not a multiple of num_kv_heads %d", num_heads_out, n_kv);
    PGK_REQUIRE(pos_buf || (pos >= 0 && pos + seq <= max_seq), "kv_cache_update failed: position %d + %d exceeds max_seq_len %d", pos, seq, max_seq);
    size_t rb = (size_t)head_dim * elem_bytes;
    size_t rows = (size_t)seq * num_heads_out;
    if (rows == 0) return 0;
    bool vec = (rb % 16 == 0) && ((uintptr_t)new_kv % 16 == 0) && ((uintptr_t)cache % 16 == 0);
    if (vec) {
        size_t n = rows * (rb / 16);
        PGK_CUDA(launch(kv_update_kernel, dim3(blocks_for(n, 256) > 4096 ? 4096 : blocks_for(n, 256)), dim3(256), 0, ST(stream),
                        (const uint8_t*)new_kv, (uint8_t*)cache, seq, n_kv, num_heads_out, rb, max_seq, pos, pos_buf));
    } else {
        size_t n = rows * rb;
        PGK_CUDA(launch(kv_update_bytes_kernel, dim3(blocks_for(n, 256) > 4096 ? 4096 : blocks_for(n, 256)), dim3(256), 0, ST(stream),
                        (const uint8_t*)new_kv, (uint8_t*)cache, seq, n_kv, num_heads_out, rb, max_seq, pos, pos_buf));
    }
    return 0;
}
int pgk_qk_norm_rope_kv_write(void* qkv, const void* q_gamma, const void* k_gamma, float eps, const float* cos_table,
                              const float* sin_table, void* k_cache, void* v_cache, int M, int n_heads_q, int n_kv,
                              int head_dim, int max_seq, int pos, const int32_t* pos_buf, void* stream) {
    PGK_REQUIRE(head_dim % 2 == 0 && M >= 1, "qk_norm_rope_kv_write failed: invalid shape");
    long warps = (long)M * (n_heads_q + 2 * n_kv);
    PGK_CUDA(launch(qk_norm_rope_kv_kernel, dim3(blocks_for((size_t)warps * 32, 128)), dim3(128), 0, ST(stream),
                    (uint16_t*)qkv, (const uint16_t*)q_gamma, (const uint16_t*)k_gamma, eps, cos_table, sin_table,
                    (uint16_t*)k_cache, (uint16_t*)v_cache, M, n_heads_q, n_kv, head_dim, max_seq, pos, pos_buf));
    return 0;
}

}  // extern "C"
