// gemv_nvf4.cu -- W4A16 GEMV (PyGPUkit "NVF4": E2M1 nibbles, UE4M3 scale per 32-k block)
// for decode batch 1..8 on B200 (sm_100a), CUDA cores only.
//
//   C[M,N] = alpha * A[M,K] . dequant(B)      B_data u8 [K/2, N] (low nibble = even k),
//                                             B_scale u8 [ceil(K/32), N]
//
// replaces gemv_nvf4_bf16_kernel(_unrolled) (native/ops/matmul/gemv/w4a16_bf16/sm120/
// nvf4_kernels.cu:19-233), which walks K serially with one thread per column.
//
// Design (see DESIGN.md "gemv_nvf4"):
//   * the reference layout is N-contiguous, so lanes own COLS consecutive columns (one
//     128-bit / 64-bit load = COLS bytes = COLS columns x 2 k) and K is split three ways:
//     over the 16 byte-rows of a scale block (in flight together), over the warps of a CTA
//     and over KS CTAs per 512-column group.  Partial sums meet in shared memory, then in
//     an f32 workspace that the last-arriving CTA of the group folds in fixed order
//     (deterministic, no float atomics).
//   * dequantisation in registers: F2FP.F16.E2M1.UNPACK_B turns one byte into the exact f16
//     pair (even k, odd k); FHFMA (fma.rn.f32.f16) multiplies by the staged f16 activation
//     with an exact product and f32 accumulation: 1.5 instructions per weight.
//   * the UE4M3 block scale is decoded with one shift-add and applied once per 32-k block
//     (E2M1 x UE4M3 is exact in f32, so this equals the reference's pre-scaled LUT).
#include "common.cuh"
#include "gemv_common.cuh"
#include "../../include/pgk_b200.h"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace pgk {

template <int COLS> struct ColVec;
template <> struct ColVec<16> {
    uint32_t w[4];
    __device__ __forceinline__ void load(const uint8_t* p) {
        uint4 v = ldg_stream(p);
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    }
};
template <> struct ColVec<8> {
    uint32_t w[2];
    __device__ __forceinline__ void load(const uint8_t* p) {
        asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p));
    }
};

// accumulate one byte-row: COLS bytes (COLS columns x (even k, odd k)) against a2 = (a_even, a_odd)
template <int M, int COLS>
__device__ __forceinline__ void nvf4_row_fma(float (&part)[M][COLS], const uint32_t* w,
                                             const uint32_t (&a2)[M]) {
#pragma unroll
    for (int q = 0; q < COLS / 4; ++q) {
        uint32_t h0 = e2m1x2_to_f16x2<0>(w[q]);
        uint32_t h1 = e2m1x2_to_f16x2<1>(w[q]);
        uint32_t h2 = e2m1x2_to_f16x2<2>(w[q]);
        uint32_t h3 = e2m1x2_to_f16x2<3>(w[q]);
#pragma unroll
        for (int m = 0; m < M; ++m) {
            fhfma_f16_lo(part[m][4 * q + 0], h0, a2[m]); fhfma_f16_hi(part[m][4 * q + 0], h0, a2[m]);
            fhfma_f16_lo(part[m][4 * q + 1], h1, a2[m]); fhfma_f16_hi(part[m][4 * q + 1], h1, a2[m]);
            fhfma_f16_lo(part[m][4 * q + 2], h2, a2[m]); fhfma_f16_hi(part[m][4 * q + 2], h2, a2[m]);
            fhfma_f16_lo(part[m][4 * q + 3], h3, a2[m]); fhfma_f16_hi(part[m][4 * q + 3], h3, a2[m]);
        }
    }
}

struct Nvf4Params {
    const uint16_t* A;
    const uint8_t* Bd;
    const uint8_t* Bs;
    uint16_t* C;
    int K, N;            // N = number of weight columns (2*I for the fused gate/up layout)
    int KS;              // CTAs per column group
    int bpc;             // scale blocks per CTA
    bool rt_swiglu;      // run-time SwiGLU epilogue in a non-SWIGLU instantiation
    float alpha;
    float* ws;           // [G*KS][M][32*COLS] f32 partials
    unsigned* counters;  // [G], zero between launches
    GemvEpilogue epi;
};

constexpr int kNvf4Hdr = 1024;  // smem header: inv_s[8] floats, reduction scratch

// SWIGLU: columns are gate/up interleaved in groups of 8 ([g0..g7,u0..u7,g8..]); the output
// has N/2 columns: C = bf16( bf16(silu(bf16(g))) * bf16(u) ).
template <int M, int COLS, bool SWIGLU, bool XCHG = false>    // XCHG: tensor-parallel form with the all-reduce inside
__global__ void __launch_bounds__(512) gemv_nvf4_kernel(Nvf4Params p) {
    constexpr int CW = 32 * COLS;  // columns per warp / per CTA
    constexpr int RH = 8;          // byte-rows in flight per lane
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* inv_s = reinterpret_cast<float*>(smem_raw);
    uint32_t* red_u = reinterpret_cast<uint32_t*>(smem_raw + 64);
    __half* a_s = reinterpret_cast<__half*>(smem_raw + kNvf4Hdr);
    float* red_f = reinterpret_cast<float*>(smem_raw + kNvf4Hdr);  // aliases a_s after the main loop
    __shared__ unsigned s_ticket;

    const int K = p.K, N = p.N;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    const int g = blockIdx.x / p.KS, ks = blockIdx.x % p.KS;
    const int KB = K >> 5;
    const int kb0 = ks * p.bpc, kb1 = min(KB, kb0 + p.bpc);
    const int n0 = g * CW + lane * COLS;
    const bool col_ok = n0 < N;

    // first half-block of weights does not depend on the previous kernel
    ColVec<COLS> d[RH];
    int b = kb0 + warp;
    if (col_ok && b < kb1) {
#pragma unroll
        for (int r = 0; r < RH; ++r) d[r].load(p.Bd + (size_t)(b * 16 + r) * N + n0);
    }
    griddep_wait();
    griddep_launch();
    const bool swiglu = SWIGLU || p.rt_swiglu;      // rt_swiglu: the plain instantiation serving the SwiGLU layout (one hot kernel per step)
    const bool xchg = XCHG && M == 1 && !swiglu && p.epi.xchg.world > 1;
    const unsigned xseq = xchg ? *p.epi.xchg.epoch + 1u : 0u;
    unsigned xticket = 0;
    if (xchg && threadIdx.x == 0) xticket = xchg_arrive(p.epi.xchg);
    stage_activations_f16<M, false>(p.A, K, a_s, inv_s, red_u, p.epi.norm_gamma, p.epi.norm_eps);

    float acc[M][COLS];
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
        for (int j = 0; j < COLS; ++j) acc[m][j] = 0.f;

    const uint32_t* a32 = reinterpret_cast<const uint32_t*>(a_s);
    bool preloaded = true;
    for (; b < kb1; b += wpc) {
        if (col_ok) {
            float part[M][COLS];
#pragma unroll
            for (int m = 0; m < M; ++m)
#pragma unroll
                for (int j = 0; j < COLS; ++j) part[m][j] = 0.f;
            ColVec<COLS> sc;
            sc.load(p.Bs + (size_t)b * N + n0);
#pragma unroll
            for (int half = 0; half < 16 / RH; ++half) {
                if (!preloaded) {
#pragma unroll
                    for (int r = 0; r < RH; ++r)
                        d[r].load(p.Bd + (size_t)(b * 16 + half * RH + r) * N + n0);
                }
                preloaded = false;
#pragma unroll
                for (int r = 0; r < RH; ++r) {
                    uint32_t a2[M];
#pragma unroll
                    for (int m = 0; m < M; ++m) a2[m] = a32[(size_t)m * (K >> 1) + b * 16 + half * RH + r];
                    nvf4_row_fma<M, COLS>(part, d[r].w, a2);
                }
            }
#pragma unroll
            for (int j = 0; j < COLS; ++j) {
                float s = ue4m3_to_f32((sc.w[j >> 2] >> (8 * (j & 3))) & 0xffu);
#pragma unroll
                for (int m = 0; m < M; ++m) acc[m][j] = fmaf(part[m][j], s, acc[m][j]);
            }
        }
        preloaded = false;
    }

    // ---- CTA reduction over warps (a_s is dead from here on)
    __syncthreads();
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
        for (int j = 0; j < COLS; ++j)
            red_f[((size_t)warp * M + m) * CW + lane * COLS + j] = acc[m][j];
    __syncthreads();

    const bool single = (p.KS == 1);
    float* my_ws = p.ws + (size_t)blockIdx.x * M * CW;
    if (!single) {
        for (int t = threadIdx.x; t < M * CW; t += blockDim.x) {
            float v = 0.f;
            for (int w = 0; w < wpc; ++w) v += red_f[(size_t)w * M * CW + t];
            __stcg(my_ws + t, v);
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_ticket = atomicAdd(p.counters + g, 1u);
        __syncthreads();
        if (s_ticket != (unsigned)(p.KS - 1)) {
            if (xchg && threadIdx.x == 0) xchg_finish(p.epi.xchg, xseq, xticket, gridDim.x);
            return;
        }
        __threadfence();
    }
    // ---- final fold (this CTA is the only / the last one of its column group)
    const float* grp_ws = p.ws + (size_t)g * p.KS * M * CW;
    auto total = [&](int m, int col) -> float {
        float v = 0.f;
        if (single) {
            for (int w = 0; w < wpc; ++w) v += red_f[((size_t)w * M + m) * CW + col];
        } else {
            for (int s0 = 0; s0 < p.KS; s0 += 8) {           // eight partials requested together, added in split order
                float t[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    t[u] = (s0 + u < p.KS) ? __ldcg(grp_ws + ((size_t)(s0 + u) * M + m) * CW + col) : 0.f;
#pragma unroll
                for (int u = 0; u < 8; ++u) v += t[u];
            }
        }
        return v * inv_s[m];
    };
    if (!swiglu) {
        for (int t = threadIdx.x; t < M * CW; t += blockDim.x) {
            int m = t / CW, col = t % CW, n = g * CW + col;
            if (n >= N) continue;
            float v = p.alpha * total(m, col);
            size_t o = (size_t)m * N + n;
            if (xchg) {                // row-parallel layer under tensor parallelism: all-reduce, then the residual
                p.C[o] = xchg_reduce_1t(p.epi.xchg, xseq, n, f32_to_bf16_bits(v), p.epi.residual);
                continue;
            }
            if (p.epi.bias) v = round_bf16(v) + bf16_bits_to_f32(p.epi.bias[n]);
            p.C[o] = p.epi.residual ? hadd_bf16_bits(v, p.epi.residual[o]) : f32_to_bf16_bits(v);
        }
        if (xchg && threadIdx.x == 0) xchg_finish(p.epi.xchg, xseq, xticket, gridDim.x);
    } else {
        const int I = N >> 1;
        for (int t = threadIdx.x; t < M * (CW / 2); t += blockDim.x) {
            int m = t / (CW / 2), j = t % (CW / 2);
            int cg = 16 * (j >> 3) + (j & 7);
            int i = g * (CW / 2) + j;
            if (i >= I) continue;
            float gate = round_bf16(p.alpha * total(m, cg));
            float up = round_bf16(p.alpha * total(m, cg + 8));
            float s = round_bf16(silu_ref(gate));
            p.C[(size_t)m * I + i] = f32_to_bf16_bits(s * up);
        }
    }
    if (!single && threadIdx.x == 0) p.counters[g] = 0;
}

// Generic fallback (any K / N / alignment): one thread per (m, column), serial over K in the
// reference's own order (nvf4_kernels.cu:19-73).
__global__ void gemv_nvf4_generic_kernel(const uint16_t* __restrict__ A, const uint8_t* __restrict__ Bd,
                                         const uint8_t* __restrict__ Bs, uint16_t* __restrict__ C, int M,
                                         int K, int N, float alpha, GemvEpilogue epi) {
    griddep_wait();
    griddep_launch();
    long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long)M * N) return;
    int n = (int)(t % N), m = (int)(t / N);
    float acc = 0.f;
    for (int k = 0; k < K; k += 2) {
        float s = ue4m3_to_f32(Bs[(size_t)(k >> 5) * N + n]);
        uint32_t byte = Bd[(size_t)(k >> 1) * N + n];
        acc = fmaf(bf16_bits_to_f32(A[(size_t)m * K + k]), e2m1_to_f32(byte & 0xf) * s, acc);
        if (k + 1 < K)
            acc = fmaf(bf16_bits_to_f32(A[(size_t)m * K + k + 1]), e2m1_to_f32(byte >> 4) * s, acc);
    }
    float v = alpha * acc;
    size_t o = (size_t)m * N + n;
    if (epi.bias) v = round_bf16(v) + bf16_bits_to_f32(epi.bias[n]);
    C[o] = epi.residual ? hadd_bf16_bits(v, epi.residual[o]) : f32_to_bf16_bits(v);
}

// ---------------------------------------------------------------- host side
static float* g_ws = nullptr;
static unsigned* g_counters = nullptr;
static const size_t kWsBytes = 48ull << 20;
static const int kMaxGroups = 4096;

static int ensure_workspace(cudaStream_t st) {
    if (g_ws) return 0;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cs);
    PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                "gemv_nvf4_bf16 failed: first call must happen outside graph capture (workspace allocation)");
    PGK_CUDA(cudaMalloc(&g_ws, kWsBytes));
    PGK_CUDA(cudaMalloc(&g_counters, kMaxGroups * sizeof(unsigned)));
    PGK_CUDA(cudaMemset(g_counters, 0, kMaxGroups * sizeof(unsigned)));
    return 0;
}

struct Nvf4Cfg { int wpc, KS, bpc; size_t smem; };

template <int M, int COLS, bool SWIGLU, bool XCHG = false>
static int launch_nvf4(Nvf4Params p, cudaStream_t st) {
    constexpr int CW = 32 * COLS;
    auto kern = gemv_nvf4_kernel<M, COLS, SWIGLU, XCHG>;
    static std::mutex mu;
    static std::unordered_map<uint64_t, Nvf4Cfg> cache;
    static bool attr_set = false;
    const int G = (p.N + CW - 1) / CW, KB = p.K >> 5;
    Nvf4Cfg cfg;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!attr_set) {
            PGK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
            attr_set = true;
        }
        uint64_t key = ((uint64_t)(uint32_t)p.K << 32) | (uint32_t)p.N;
        auto it = cache.find(key);
        if (it == cache.end()) {
            const int sms = sm_count();
            Nvf4Cfg best = {8, 1, KB, 0};
            double best_score = -1.0;
            for (int wpc = 2; wpc <= 16; ++wpc) {
                size_t body = (size_t)M * p.K * 2;
                size_t redb = (size_t)wpc * M * CW * 4;
                size_t smem = kNvf4Hdr + (body > redb ? body : redb);
                if (smem > 227 * 1024) continue;
                int occ = 0;
                PGK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, wpc * 32, smem));
                for (int c = 1; c <= occ; ++c) {
                    long T = (long)sms * c;
                    int KS = (int)(T / G);
                    if (KS < 1) KS = 1;
                    if (KS > KB) KS = KB;
                    int bpc = (KB + KS - 1) / KS;
                    KS = (KB + bpc - 1) / bpc;
                    int bw = (bpc + wpc - 1) / wpc;
                    long ctas = (long)G * KS;
                    long per_sm = (ctas + sms - 1) / sms;
                    double eff = (double)G * KB / ((double)sms * per_sm * wpc * bw);
                    int W = (int)(per_sm * wpc);
                    double score = eff + 0.0004 * (W > 32 ? 32 : W) - 0.0002 * c;
                    if ((size_t)ctas * M * CW * 4 > kWsBytes || G > kMaxGroups) continue;
                    if (score > best_score) { best_score = score; best = {wpc, KS, bpc, smem}; }
                }
            }
            PGK_REQUIRE(best_score > 0, "gemv_nvf4_bf16 failed: no launch configuration (K=%d N=%d M=%d)", p.K, p.N, M);
            // Few column groups (N <= 74 x CW, e.g. the 8192-wide o_proj / down_proj of Llama-3-70B): measured on B200
            // (scripts/sweep_nvf4.py, K = N = 8192) one fat CTA per SM -- 16 warps, floor(SMs / groups) K-splits, at
            // least 8 scale blocks per CTA -- beats the many-thin-CTA choice of the score above by 2.5x (12.0 vs 29.9 us):
            // fewer activation stagings and partial folds, same bytes in flight.
            if (2 * G <= sms && G <= kMaxGroups) {
                int KS = sms / G;
                if (KS > KB / 8) KS = KB / 8 > 1 ? KB / 8 : 1;
                int bpc = (KB + KS - 1) / KS;
                KS = (KB + bpc - 1) / bpc;
                int wpc = bpc < 16 ? (bpc < 4 ? 4 : bpc) : 16;
                size_t body = (size_t)M * p.K * 2, redb = (size_t)wpc * M * CW * 4;
                size_t smem = kNvf4Hdr + (body > redb ? body : redb);
                int occ = 0;
                if (smem <= 227 * 1024 && (size_t)G * KS * M * CW * 4 <= kWsBytes &&
                    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, wpc * 32, smem) == cudaSuccess && occ >= 1)
                    best = {wpc, KS, bpc, smem};
            }
            if (const char* ov = getenv("PGK_NVF4_CFG")) {       // tuning aid: "warps_per_cta,k_splits"
                int w = 0, ks = 0;
                if (sscanf(ov, "%d,%d", &w, &ks) == 2 && w >= 2 && w <= 16 && ks >= 1 && ks <= KB) {
                    int bpc = (KB + ks - 1) / ks;
                    ks = (KB + bpc - 1) / bpc;
                    size_t body = (size_t)M * p.K * 2, redb = (size_t)w * M * CW * 4;
                    best = {w, ks, bpc, kNvf4Hdr + (body > redb ? body : redb)};
                }
            }
            it = cache.emplace(key, best).first;
        }
        cfg = it->second;
    }
    p.KS = cfg.KS;
    p.bpc = cfg.bpc;
    p.ws = g_ws;
    p.counters = g_counters;
    PGK_CUDA(launch(kern, dim3(G * cfg.KS), dim3(cfg.wpc * 32), cfg.smem, st, p));
    return 0;
}

// Under tensor parallelism the row-parallel layers run the XCHG instantiation; measured at 2 GPUs (Llama-3-70B NVF4): when
// the column-parallel GEMVs of the same step run the plain instantiation, the step alternates between two 41 KB kernels
// and loses 2.2 us per launch to instruction fetch (132.7 vs 139.9 tok/s) -- so with peers open every M = 1 GEMV runs the
// XCHG code (its exchange switched off by world = 0) and the step keeps one hot kernel.
static bool tp_active() {
    GemvXchg x;
    return p2p_gemv_xchg(&x) && x.world > 1;
}

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// gemm_tc_nvf4.cu: tcgen05 skinny GEMM (exact BF16 dequantisation, weights read once for up to 64 rows)
bool gemm_nvf4_tc_shape_ok(int M, int K, int N, const void* A, const void* Bd, const void* Bs);
int gemm_nvf4_tc(const void* A, const void* Bd, const void* Bs, void* C, int M, int K, int N, float alpha, const void* bias,
                 const void* residual, cudaStream_t st);
int gemm_fp8_tc_min_m();    // gemm_tc.cu: smallest batch sent to the tensor cores (PGK_TC_MIN_M, default 5)

static int nvf4_dispatch(const void* A, const void* Bd, const void* Bs, void* C, int M, int K, int N,
                         float alpha, const void* bias, const void* residual, bool swiglu,
                         cudaStream_t st, const void* gamma = nullptr, float eps = 0.f) {
    PGK_REQUIRE(M >= 1 && K >= 1 && N >= 1, "gemv_nvf4_bf16 failed: invalid shape M=%d K=%d N=%d", M, K, N);
    const bool fast = (K % 32 == 0) && (N % 16 == 0) && al16(A) && al16(Bd) && al16(Bs) &&
                      ((size_t)K * 2 + kNvf4Hdr <= 200 * 1024);
    PGK_REQUIRE(fast || !swiglu, "gemv_nvf4_swiglu_bf16 failed: needs K%%32==0, N%%16==0 and 16-byte aligned pointers");
    PGK_REQUIRE(!gamma || (fast && K >= 512 && al16(gamma)),
                "rmsnorm_gemv_nvf4_bf16 failed: the fused norm needs the fast path and K >= 512 (K=%d)", K);
    const int Nout = swiglu ? N / 2 : N;
    if (!fast) {
        GemvEpilogue epi = {(const uint16_t*)bias, (const uint16_t*)residual};
        long tot = (long)M * N;
        PGK_CUDA(launch(gemv_nvf4_generic_kernel, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st,
                        (const uint16_t*)A, (const uint8_t*)Bd, (const uint8_t*)Bs, (uint16_t*)C, M, K, N,
                        alpha, epi));
        return 0;
    }
    // decode batch >= 5: the tensor-core GEMM reads the weights once for all rows (PGK_TC=0 keeps the slabs of four rows)
    static const bool tc_on = !(getenv("PGK_TC") && atoi(getenv("PGK_TC")) == 0);
    // measured crossover (profiles/r02_batch_nvf4_*.txt): five rows are still faster as 4 + 1 on the CUDA cores
    if (tc_on && !swiglu && !gamma && M >= gemm_fp8_tc_min_m() + 1 && gemm_nvf4_tc_shape_ok(M, K, N, A, Bd, Bs))
        return gemm_nvf4_tc(A, Bd, Bs, C, M, K, N, alpha, bias, residual, st);
    if (int rc = ensure_workspace(st)) return rc;
    int max_m = 4;
    while (max_m > 1 && (size_t)max_m * K * 2 + kNvf4Hdr > 200 * 1024) max_m >>= 1;
    for (int m0 = 0; m0 < M;) {
        int mt = M - m0 > max_m ? max_m : M - m0;
        if (mt == 3) mt = 2;
        Nvf4Params p = {};
        p.A = (const uint16_t*)A + (size_t)m0 * K;
        p.Bd = (const uint8_t*)Bd;
        p.Bs = (const uint8_t*)Bs;
        p.C = (uint16_t*)C + (size_t)m0 * Nout;
        p.K = K; p.N = N; p.alpha = alpha;
        p.epi = {(const uint16_t*)bias, residual ? (const uint16_t*)residual + (size_t)m0 * N : nullptr,
                 (const uint16_t*)gamma, eps};
        int rc;
        if (swiglu) {
            static const bool one_kernel = !(getenv("PGK_NVF4_ONE_KERNEL") && atoi(getenv("PGK_NVF4_ONE_KERNEL")) == 0);
            if (mt == 1 && one_kernel) {
                // the gate/up launch runs the SAME instantiation as qkv / o / down (SwiGLU epilogue chosen at run time):
                // a decode step that alternates between two 41 KB GEMV kernels pays for the instruction fetch
                p.rt_swiglu = true;
                rc = tp_active() ? launch_nvf4<1, 16, false, true>(p, st) : launch_nvf4<1, 16, false>(p, st);
            }
            else if (mt == 1) rc = launch_nvf4<1, 16, true>(p, st);
            else if (mt == 2) rc = launch_nvf4<2, 16, true>(p, st);
            else              rc = launch_nvf4<4, 8, true>(p, st);
        } else {
            if (mt == 1 && tp_active())  rc = launch_nvf4<1, 16, false, true>(p, st);     // see tp_active()
            else if (mt == 1) rc = launch_nvf4<1, 16, false>(p, st);
            else if (mt == 2) rc = launch_nvf4<2, 16, false>(p, st);
            else              rc = launch_nvf4<4, 8, false>(p, st);
        }
        if (rc) return rc;
        m0 += mt;
    }
    return 0;
}

// ---------------------------------------------------------------- allocation-free variant (pgk_gemv_nvf4_bf16_noalloc)
// The split-K partials and the per-column-group arrival counters of gemv_nvf4_kernel live in a file-scope device array
// instead of the cudaMalloc'ed workspace above: 1 MB of f32 partials (<= 512 CTAs x 512 columns) + 1024 counters.  Every
// GPU that loads the library holds these 1028 KB until exit.  All calls share the array: calls that overlap on the GPU
// (different streams, or two processes' worth of work on one context) race on it.
constexpr int kNaMaxCtas = 512;                      // CTAs that write partials (column groups x K splits)
constexpr int kNaMaxGroups = 1024;                   // column groups of 512 columns: N <= 524288
constexpr size_t kNaWsFloats = (size_t)kNaMaxCtas * 512;
static __device__ __align__(256) unsigned char g_na_scratch[kNaWsFloats * 4 + kNaMaxGroups * 4];

// The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that
// device current.
static int na_scratch_ptr(unsigned char** out) {
    static std::mutex mu;
    static unsigned char* ptrs[64] = {};
    int dev = 0;
    PGK_CUDA(cudaGetDevice(&dev));
    PGK_REQUIRE(dev >= 0 && dev < 64, "gemv_nvf4_bf16_noalloc failed: device ordinal %d", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[dev]) {
        void* ptr = nullptr;
        PGK_CUDA(cudaGetSymbolAddress(&ptr, g_na_scratch));
        PGK_REQUIRE(ptr != nullptr, "gemv_nvf4_bf16_noalloc failed: cudaGetSymbolAddress returned a null address");
        ptrs[dev] = (unsigned char*)ptr;
    }
    *out = ptrs[dev];
    return 0;
}

static int launch_nvf4_noalloc(Nvf4Params p, cudaStream_t st) {
    constexpr int CW = 512;
    auto kern = gemv_nvf4_kernel<1, 16, false>;
    const int G = (p.N + CW - 1) / CW, KB = p.K >> 5;
    PGK_REQUIRE(G <= kNaMaxGroups, "gemv_nvf4_bf16_noalloc failed: N=%d exceeds the static scratch (N <= %d)", p.N, kNaMaxGroups * CW);
    unsigned char* scratch = nullptr;
    if (int rc = na_scratch_ptr(&scratch)) return rc;
    // one fat CTA per SM where the column groups are few (the choice launch_nvf4 measured best), never more partial
    // records than the scratch holds; more than 512 column groups run unsplit and use no partials at all
    int KS = sm_count() / G;
    if (KS > KB / 8) KS = KB / 8;
    if (KS > kNaMaxCtas / G) KS = kNaMaxCtas / G;
    if (KS < 1) KS = 1;
    const int bpc = (KB + KS - 1) / KS;
    KS = (KB + bpc - 1) / bpc;
    const int wpc = bpc < 16 ? (bpc < 4 ? 4 : bpc) : 16;
    const size_t body = (size_t)p.K * 2, redb = (size_t)wpc * CW * 4;
    const size_t smem = kNvf4Hdr + (body > redb ? body : redb);
    PGK_REQUIRE(smem <= 226 * 1024, "gemv_nvf4_bf16_noalloc failed: K=%d too large", p.K);
    PGK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    p.KS = KS;
    p.bpc = bpc;
    p.ws = reinterpret_cast<float*>(scratch);
    p.counters = reinterpret_cast<unsigned*>(scratch + kNaWsFloats * 4);
    // the array keeps the last call's contents: the counters must start at zero (the partials are overwritten before
    // they are read)
    if (KS > 1) PGK_CUDA(cudaMemsetAsync(p.counters, 0, (size_t)G * sizeof(unsigned), st));
    PGK_CUDA(launch(kern, dim3(G * KS), dim3(wpc * 32), smem, st, p));
    return 0;
}

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_gemv_nvf4_bf16(const void* A, const void* B_data, const void* B_scale, void* C, int M, int K,
                       int N, float alpha, void* stream) {
    return nvf4_dispatch(A, B_data, B_scale, C, M, K, N, alpha, nullptr, nullptr, false, (cudaStream_t)stream);
}
int pgk_gemv_nvf4_bf16_ex(const void* A, const void* B_data, const void* B_scale, void* C, int M, int K,
                          int N, float alpha, const void* bias, const void* residual, void* stream) {
    return nvf4_dispatch(A, B_data, B_scale, C, M, K, N, alpha, bias, residual, false, (cudaStream_t)stream);
}
int pgk_gemv_nvf4_swiglu_bf16(const void* A, const void* Wgu_data, const void* Wgu_scale, void* C, int M,
                              int K, int N2, void* stream) {
    return nvf4_dispatch(A, Wgu_data, Wgu_scale, C, M, K, N2, 1.0f, nullptr, nullptr, true, (cudaStream_t)stream);
}

/* NVF4 form of pgk_gemv_fp8_allreduce_bf16: row-parallel linear of one token + all-reduce over the ranks + residual in one launch */
int pgk_gemv_nvf4_allreduce_bf16(const void* A, const void* B_data, const void* B_scale, const void* residual, void* out,
                                 int K, int N, float alpha, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    Nvf4Params p = {};
    p.epi = {nullptr, (const uint16_t*)residual};
    PGK_REQUIRE(p2p_gemv_xchg(&p.epi.xchg), "gemv_nvf4_allreduce failed: peer buffers not opened (pgk_p2p_create / pgk_p2p_open)");
    PGK_REQUIRE(K % 32 == 0 && N % 16 == 0 && al16(A) && al16(B_data) && al16(B_scale) && (size_t)K * 2 + kNvf4Hdr <= 200 * 1024 &&
                N >= 16 && (unsigned)N <= p.epi.xchg.slot_words,
                "gemv_nvf4_allreduce failed: needs K %% 32 == 0, N %% 16 == 0, N <= %u and 16-byte aligned pointers (K=%d N=%d)",
                p.epi.xchg.slot_words, K, N);
    if (p.epi.xchg.world <= 1) p.epi.xchg.world = 0;
    if (int rc = ensure_workspace(st)) return rc;
    p.A = (const uint16_t*)A; p.Bd = (const uint8_t*)B_data; p.Bs = (const uint8_t*)B_scale; p.C = (uint16_t*)out;
    p.K = K; p.N = N; p.alpha = alpha;
    return launch_nvf4<1, 16, false, true>(p, st);
}

/* pgk_gemv_nvf4_bf16 for callers that cannot allocate (CUDA graph capture, tight serving loops): no cudaMalloc / cudaFree,
 * the split-K scratch is a 1028 KB file-scope device array shared by every call (overlapping calls race).  Rows are
 * served one at a time by the M = 1 kernel.  Limits: K % 32 == 0, K <= 100 K, N % 16 == 0, N <= 524288, 16-byte aligned. */
int pgk_gemv_nvf4_bf16_noalloc(const void* A, const void* B_data, const void* B_scale, void* C, int M, int K, int N,
                               float alpha, void* stream) {
    PGK_REQUIRE(M >= 1 && K >= 32 && N >= 16, "gemv_nvf4_bf16_noalloc failed: invalid shape M=%d K=%d N=%d", M, K, N);
    PGK_REQUIRE(K % 32 == 0 && N % 16 == 0 && al16(A) && al16(B_data) && al16(B_scale) && (size_t)K * 2 + kNvf4Hdr <= 200 * 1024,
                "gemv_nvf4_bf16_noalloc failed: needs K %% 32 == 0, K <= 101888, N %% 16 == 0 and 16-byte aligned pointers (K=%d N=%d)", K, N);
    PGK_REQUIRE(N <= kNaMaxGroups * 512, "gemv_nvf4_bf16_noalloc failed: N=%d exceeds the static scratch (N <= %d)", N, kNaMaxGroups * 512);
    for (int m = 0; m < M; ++m) {
        Nvf4Params p = {};
        p.A = (const uint16_t*)A + (size_t)m * K; p.Bd = (const uint8_t*)B_data; p.Bs = (const uint8_t*)B_scale;
        p.C = (uint16_t*)C + (size_t)m * N;
        p.K = K; p.N = N; p.alpha = alpha;
        if (int rc = launch_nvf4_noalloc(p, (cudaStream_t)stream)) return rc;
    }
    return 0;
}

/* rmsnorm(x, gamma, eps) -> NVF4 GEMV in one launch (see pgk_rmsnorm_gemv_fp8_bf16). */
int pgk_rmsnorm_gemv_nvf4_bf16(const void* A, const void* gamma, float eps, const void* B_data,
                               const void* B_scale, void* C, int M, int K, int N, float alpha, const void* bias,
                               const void* residual, void* stream) {
    return nvf4_dispatch(A, B_data, B_scale, C, M, K, N, alpha, bias, residual, false, (cudaStream_t)stream,
                         gamma, eps);
}
int pgk_rmsnorm_gemv_nvf4_swiglu_bf16(const void* A, const void* gamma, float eps, const void* Wgu_data,
                                      const void* Wgu_scale, void* C, int M, int K, int N2, void* stream) {
    return nvf4_dispatch(A, Wgu_data, Wgu_scale, C, M, K, N2, 1.0f, nullptr, nullptr, true,
                         (cudaStream_t)stream, gamma, eps);
}

}  // extern "C"
