// gemv_common.cuh -- pieces shared by the weight-streaming GEMV kernels
// (gemv_fp8.cu, gemv_nvf4.cu, gemv_bf16.cu).
#pragma once
#include "common.cuh"

namespace pgk {

struct GemvCfg {
    int warps_per_cta;
    int ctas;
    int rho;  // rows per warp
};

// Pick (warps per CTA, CTAs) such that all CTAs are co-resident (one wave) and the rows
// divide evenly over the warps.  max_ctas_per_sm[wpc] comes from the occupancy calculator.
inline GemvCfg pick_gemv_cfg(int n_rows, const int* max_ctas_per_sm /* index = warps per CTA, 1..16 */) {
    const int sms = sm_count();
    GemvCfg best = {8, sms, (n_rows + sms * 8 - 1) / (sms * 8)};
    double best_score = -1.0;
    for (int wpc = 2; wpc <= 16; ++wpc) {
        for (int c = 1; c <= max_ctas_per_sm[wpc]; ++c) {
            int W = wpc * c;
            long total = (long)sms * W;
            int rho = (int)((n_rows + total - 1) / total);
            double eff = (double)n_rows / ((double)rho * (double)total);
            // prefer high efficiency; tie-break toward more resident warps (bytes in flight)
            // and fewer, larger CTAs (less activation staging)
            double score = eff + 0.0004 * (W > 40 ? 40 : W) - 0.0002 * c;
            if (score > best_score) {
                best_score = score;
                best = {wpc, sms * c, rho};
            }
        }
    }
    return best;
}

struct GemvEpilogue {
    const uint16_t* bias;      // [N] bf16 or null
    const uint16_t* residual;  // [M,N] bf16 or null
    const uint16_t* norm_gamma = nullptr;  // [K] bf16: A <- bf16(rmsnorm(A) * gamma) while staging, or null
    float norm_eps = 0.f;
    GemvXchg xchg = {};                    // world > 1: all-reduce over the ranks before the residual add (M == 1)
};

// ---- in-kernel all-reduce of one output element (see GemvXchg).  `lanes` cooperating threads (lane index `l`): lane r
// publishes to rank r and fetches rank r's word; the caller sums `vals` in rank order.  Returns this lane's fetched value
// (its own partial for r == rank); lanes >= world return 0.
__device__ __forceinline__ void xchg_publish(const GemvXchg& x, unsigned seq, int n, uint16_t mine, int r) {
    unsigned char* dst = x.ll[r] + (size_t)(seq & 1u) * x.par_bytes + ((size_t)x.rank * x.slot_words + (size_t)n) * 8;
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(dst), "r"((uint32_t)mine), "r"(seq) : "memory");
}
__device__ __forceinline__ float xchg_fetch(const GemvXchg& x, unsigned seq, int n, int r) {
    const unsigned char* src = x.ll[x.rank] + (size_t)(seq & 1u) * x.par_bytes + ((size_t)r * x.slot_words + (size_t)n) * 8;
    uint32_t d, f;
    const long long t0 = clock64();
    for (;;) {
        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d), "=r"(f) : "l"(src) : "memory");
        if (f == seq) break;
        if (clock64() - t0 > 4000000000LL) { *x.err = 1u; break; }        // ~2 s: a peer is gone; do not hang the GPU
    }
    return bf16_bits_to_f32(d & 0xffffu);
}
// one thread does the whole element (used where a thread, not a warp, owns an output column)
__device__ __forceinline__ uint16_t xchg_reduce_1t(const GemvXchg& x, unsigned seq, int n, uint16_t mine, const uint16_t* residual) {
    for (int r = 0; r < x.world; ++r)
        if (r != x.rank) xchg_publish(x, seq, n, mine, r);
    float acc = 0.f;
    for (int r = 0; r < x.world; ++r) acc += (r == x.rank) ? bf16_bits_to_f32(mine) : xchg_fetch(x, seq, n, r);
    return residual ? hadd_bf16_bits(acc, residual[n]) : f32_to_bf16_bits(acc);
}
// Advancing the exchange number: every CTA reads `epoch` after the dependency wait and then counts itself in
// (xchg_arrive, one atomic whose result is not needed until the CTA ends, so its latency hides behind the GEMV); the CTA
// that was the last to read stores the new number when it finishes (xchg_finish) -- no CTA can still need the old one.
__device__ __forceinline__ unsigned xchg_arrive(const GemvXchg& x) { return atomicAdd(x.done, 1u); }
__device__ __forceinline__ void xchg_finish(const GemvXchg& x, unsigned seq, unsigned ticket, unsigned participants) {
    if (ticket == participants - 1) { *x.done = 0; *x.epoch = seq; }
}

// ---- fused RMSNorm prologue -------------------------------------------------------------------
// inv_rms of one activation row in EXACTLY the summation order of the stand-alone rmsnorm kernel
// (= the reference's rmsnorm_bf16_kernel, norm_kernels.cuh:535-585): bs = min(256, K^32) virtual
// threads, thread t folds elements t, t+bs, ... with fma, a shuffle tree per warp, then a tree over
// the warp totals.  `scratch` needs bs floats of shared memory; every thread returns the same value,
// so fused and unfused paths give bit-identical normalised activations.
__device__ __forceinline__ float row_inv_rms_ref_order(const uint16_t* __restrict__ row, int K, float eps,
                                                       float* scratch) {
    const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31;
    const int bs = min(256, (K + 31) & ~31), nw = bs >> 5;
    for (int v = tid; v < bs; v += nthr) {
        // the first 16 elements of the virtual thread are loaded together (16 loads in flight), the
        // fma chain then runs in index order; padding zeros leave the partial sum unchanged
        float xv[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            const int i = v + e * bs;
            xv[e] = (i < K) ? bf16_bits_to_f32(ld_act2(row + i)) : 0.f;
        }
        float part = 0.f;
#pragma unroll
        for (int e = 0; e < 16; ++e) part = fmaf(xv[e], xv[e], part);
        for (int i = v + 16 * bs; i < K; i += bs) { float x = bf16_bits_to_f32(ld_act2(row + i)); part = fmaf(x, x, part); }
        scratch[v] = part;
    }
    __syncthreads();
    float tv = 0.f;
    for (int w = 0; w < nw; ++w) {
        float t = warp_sum(scratch[w * 32 + lane]);
        if (lane == w) tv = t;
    }
    const float ss = warp_sum(tv);
    __syncthreads();                       // scratch is free again
    return rsqrtf(ss / (float)K + eps);
}
// eight elements: bf16( x * inv * gamma ) (left to right, as rmsnorm_bf16_kernel), packed as bf16 bits
__device__ __forceinline__ uint4 norm8_bf16(const uint4& x, const uint4& g, float inv) {
    uint4 o;
    o.x = pack_bf16x2(bf16lo(x.x) * inv * bf16lo(g.x), bf16hi(x.x) * inv * bf16hi(g.x));
    o.y = pack_bf16x2(bf16lo(x.y) * inv * bf16lo(g.y), bf16hi(x.y) * inv * bf16hi(g.y));
    o.z = pack_bf16x2(bf16lo(x.z) * inv * bf16lo(g.z), bf16hi(x.z) * inv * bf16hi(g.z));
    o.w = pack_bf16x2(bf16lo(x.w) * inv * bf16lo(g.w), bf16hi(x.w) * inv * bf16hi(g.w));
    return o;
}

// Stage A[M,K] (bf16) into shared memory as range-shifted f16.  Layout per row: four planes
// [K/32 chunks][8 halves], plane p holding k%32 in [8p, 8p+8), so that the consumer's
// per-lane 16-byte reads (lane = chunk) are bank-conflict free.
// PLANES = false keeps the natural k order (used by the column-parallel NVF4 kernel).
template <int M, bool PLANES = true>
__device__ __forceinline__ void stage_activations_f16_3pass(const uint16_t* __restrict__ A, int K,
                                                            __half* a_s, float* inv_s, uint32_t* red,
                                                            const uint16_t* __restrict__ gamma = nullptr,
                                                            float eps = 0.f) {
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int nvec = K >> 3;  // 16-byte vectors per row
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    float rinv[M];
#pragma unroll
    for (int m = 0; m < M; ++m)
        rinv[m] = gamma ? row_inv_rms_ref_order(A + (size_t)m * K, K, eps, reinterpret_cast<float*>(a_s)) : 1.f;
    const uint4* gv = reinterpret_cast<const uint4*>(gamma);
    uint32_t mx[M];
#pragma unroll
    for (int m = 0; m < M; ++m) mx[m] = 0;
#pragma unroll
    for (int m = 0; m < M; ++m) {
        const uint4* src = reinterpret_cast<const uint4*>(A + (size_t)m * K);
        for (int j = tid; j < nvec; j += nthr) {
            uint4 v = ld_act16(src + j);
            if (gamma) v = norm8_bf16(v, __ldg(gv + j), rinv[m]);
            uint32_t t = __vmaxu2(__vmaxu2(v.x & 0x7fff7fffu, v.y & 0x7fff7fffu),
                                  __vmaxu2(v.z & 0x7fff7fffu, v.w & 0x7fff7fffu));
            mx[m] = max(mx[m], max(t & 0xffffu, t >> 16));
        }
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx[m] = max(mx[m], __shfl_xor_sync(0xffffffffu, mx[m], o));
        if (lane == 0) red[m * 16 + warp] = mx[m];
    }
    __syncthreads();
    float mult[M];
#pragma unroll
    for (int m = 0; m < M; ++m) {
        uint32_t v = 0;
        for (int w = 0; w < nwarp; ++w) v = max(v, red[m * 16 + w]);
        int e = (int)(v >> 7);           // biased bf16/f32 exponent of max|a|
        e = min(max(e, 15), 254);        // keep both 2^(141-e) and 2^(e-141) normal
        if ((v >> 7) == 255) e = 141;    // inf/nan present: no rescale, let it propagate
        mult[m] = __uint_as_float((uint32_t)(268 - e) << 23);       // 2^(141-e)
        if (tid == 0) inv_s[m] = __uint_as_float((uint32_t)(e - 14) << 23);  // 2^(e-141)
    }
    const int nchunk = K >> 5;
#pragma unroll
    for (int m = 0; m < M; ++m) {
        const uint4* src = reinterpret_cast<const uint4*>(A + (size_t)m * K);
        uint4* dst = reinterpret_cast<uint4*>(a_s + (size_t)m * K);
        for (int j = tid; j < nvec; j += nthr) {
            uint4 v = ld_act16(src + j);
            if (gamma) v = norm8_bf16(v, __ldg(gv + j), rinv[m]);
            uint4 o;
            __half2 h;
            h = __floats2half2_rn(bf16lo(v.x) * mult[m], bf16hi(v.x) * mult[m]); o.x = *reinterpret_cast<uint32_t*>(&h);
            h = __floats2half2_rn(bf16lo(v.y) * mult[m], bf16hi(v.y) * mult[m]); o.y = *reinterpret_cast<uint32_t*>(&h);
            h = __floats2half2_rn(bf16lo(v.z) * mult[m], bf16hi(v.z) * mult[m]); o.z = *reinterpret_cast<uint32_t*>(&h);
            h = __floats2half2_rn(bf16lo(v.w) * mult[m], bf16hi(v.w) * mult[m]); o.w = *reinterpret_cast<uint32_t*>(&h);
            dst[PLANES ? ((j & 3) * nchunk + (j >> 2)) : j] = o;
        }
    }
    __syncthreads();
}


__device__ __forceinline__ uint4 bf16x8_to_f16x8_scaled(const uint4& v, float mult) {
    uint4 o;
    __half2 h;
    h = __floats2half2_rn(bf16lo(v.x) * mult, bf16hi(v.x) * mult); o.x = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2half2_rn(bf16lo(v.y) * mult, bf16hi(v.y) * mult); o.y = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2half2_rn(bf16lo(v.z) * mult, bf16hi(v.z) * mult); o.z = *reinterpret_cast<uint32_t*>(&h);
    h = __floats2half2_rn(bf16lo(v.w) * mult, bf16hi(v.w) * mult); o.w = *reinterpret_cast<uint32_t*>(&h);
    return o;
}

// Single-global-read staging: each thread keeps its (<= VY) 16-byte vectors of a row in registers, so the
// only L2 round trip is the first load; the RMSNorm statistics are taken from a raw copy of the row in
// shared memory (reference summation order, see row_inv_rms_ref_order), the block maximum goes through
// `red`, and the f16 planes are written last.  Same arithmetic, hence the same bits, as the 3-pass version.
//   red[0 .. 63]   : per-warp maxima (m * 16 + warp)      red[64 .. 71] : per-virtual-warp sums of squares
template <int M, bool PLANES = true>
__device__ __forceinline__ void stage_activations_f16(const uint16_t* __restrict__ A, int K,
                                                      __half* a_s, float* inv_s, uint32_t* red,
                                                      const uint16_t* __restrict__ gamma = nullptr,
                                                      float eps = 0.f) {
    constexpr int VY = 8;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int nvec = K >> 3;
    if (nvec > VY * nthr) {
        stage_activations_f16_3pass<M, PLANES>(A, K, a_s, inv_s, red, gamma, eps);
        return;
    }
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    const int nchunk = K >> 5;
    const uint4* gv = reinterpret_cast<const uint4*>(gamma);
    float* red_f = reinterpret_cast<float*>(red) + 64;
#pragma unroll 1
    for (int m = 0; m < M; ++m) {
        const uint4* src = reinterpret_cast<const uint4*>(A + (size_t)m * K);
        uint4* dst = reinterpret_cast<uint4*>(a_s + (size_t)m * K);
        uint4 v[VY], g[VY];
#pragma unroll
        for (int e = 0; e < VY; ++e) {
            const int j = tid + e * nthr;
            if (j < nvec) {
                v[e] = ld_act16(src + j);
                if (gamma) g[e] = __ldg(gv + j);
            }
        }
        if (gamma) {
            // raw row -> shared memory (natural order) for the reference-order sum of squares
#pragma unroll
            for (int e = 0; e < VY; ++e) {
                const int j = tid + e * nthr;
                if (j < nvec) dst[j] = v[e];
            }
            __syncthreads();
            const uint16_t* raw = reinterpret_cast<const uint16_t*>(dst);
            const int bs = min(256, (K + 31) & ~31), nvw = bs >> 5;
            for (int w = warp; w < nvw; w += nwarp) {           // virtual warp w = virtual threads 32w .. 32w+31
                float part = 0.f;
                for (int i = w * 32 + lane; i < K; i += bs) { float x = bf16_bits_to_f32(raw[i]); part = fmaf(x, x, part); }
                part = warp_sum(part);
                if (lane == 0) red_f[w] = part;
            }
            __syncthreads();
            const float ss = warp_sum(lane < nvw ? red_f[lane] : 0.f);
            const float rinv = rsqrtf(ss / (float)K + eps);
#pragma unroll
            for (int e = 0; e < VY; ++e)
                if (tid + e * nthr < nvec) v[e] = norm8_bf16(v[e], g[e], rinv);
        }
        uint32_t mx = 0;
#pragma unroll
        for (int e = 0; e < VY; ++e) {
            if (tid + e * nthr < nvec) {
                uint32_t t = __vmaxu2(__vmaxu2(v[e].x & 0x7fff7fffu, v[e].y & 0x7fff7fffu),
                                      __vmaxu2(v[e].z & 0x7fff7fffu, v[e].w & 0x7fff7fffu));
                mx = max(mx, max(t & 0xffffu, t >> 16));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) red[m * 16 + warp] = mx;
        __syncthreads();                 // also: every thread is done reading the raw row
        uint32_t bm = 0;
        for (int w = 0; w < nwarp; ++w) bm = max(bm, red[m * 16 + w]);
        int ex = (int)(bm >> 7);
        ex = min(max(ex, 15), 254);
        if ((bm >> 7) == 255) ex = 141;
        const float mult = __uint_as_float((uint32_t)(268 - ex) << 23);
        if (tid == 0) inv_s[m] = __uint_as_float((uint32_t)(ex - 14) << 23);
#pragma unroll
        for (int e = 0; e < VY; ++e) {
            const int j = tid + e * nthr;
            if (j < nvec) dst[PLANES ? ((j & 3) * nchunk + (j >> 2)) : j] = bf16x8_to_f16x8_scaled(v[e], mult);
        }
    }
    __syncthreads();
}

// ---- dot-product bodies shared by the per-op GEMV kernels and the one-launch step kernel (step.cu): the same
// instruction sequence, hence the same bits, on both paths
// 8 e4m3 weights (two 32-bit words) x 8 staged f16 activations (one uint4)
__device__ __forceinline__ void dot8_e4m3(float& s0, float& s1, uint32_t w0, uint32_t w1, const uint4& a) {
    uint32_t h;
    h = e4m3x2_to_f16x2_lo(w0); fhfma_f16_lo(s0, h, a.x); fhfma_f16_hi(s1, h, a.x);
    h = e4m3x2_to_f16x2_hi(w0); fhfma_f16_lo(s0, h, a.y); fhfma_f16_hi(s1, h, a.y);
    h = e4m3x2_to_f16x2_lo(w1); fhfma_f16_lo(s0, h, a.z); fhfma_f16_hi(s1, h, a.z);
    h = e4m3x2_to_f16x2_hi(w1); fhfma_f16_lo(s0, h, a.w); fhfma_f16_hi(s1, h, a.w);
}
// sum over 32 weights (one 256-bit load) x 32 staged activations (4 planes)
__device__ __forceinline__ float dot32_e4m3(const u32x8& w, const uint4* ap, int nchunk, int c) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int p = 0; p < 4; ++p) dot8_e4m3(s0, s1, w.v[2 * p], w.v[2 * p + 1], ap[p * nchunk + c]);
    return s0 + s1;
}

__device__ __forceinline__ void dot8_bf16(float& s0, float& s1, const uint32_t* w, const uint4& a) {
    fhfma_bf16_lo(s0, w[0], a.x); fhfma_bf16_hi(s1, w[0], a.x);
    fhfma_bf16_lo(s0, w[1], a.y); fhfma_bf16_hi(s1, w[1], a.y);
    fhfma_bf16_lo(s0, w[2], a.z); fhfma_bf16_hi(s1, w[2], a.z);
    fhfma_bf16_lo(s0, w[3], a.w); fhfma_bf16_hi(s1, w[3], a.w);
}

__device__ __forceinline__ unsigned long long argmax_key(float v, int idx) {
    uint32_t u = __float_as_uint(v);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);   // monotone float -> uint
    return ((unsigned long long)u << 32) | (uint32_t)(0x7fffffff - idx);
}

__device__ __forceinline__ float silu_ref(float x) {
    // silu_bf16_kernel: x / (1 + expf(-x)) in f32 (activation_kernels.cuh:40-42)
    return x / (1.0f + __expf(-x));
}



}  // namespace pgk
