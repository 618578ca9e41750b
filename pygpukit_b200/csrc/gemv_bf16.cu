// gemv_bf16.cu -- BF16 GEMV C[M,N] = A[M,K] . B[N,K]^T for lm_head / router / BF16 layers,
// with an optional fused greedy argmax (lm_head + sample_greedy in one pass).
//
// replaces gemv_bf16_opt(_vec8)_kernel (native/ops/matmul/gemv/bf16_bf16/sm120/
// bf16_opt.cuh:56-231, bf16_opt.cu:36-47) and sample_argmax_bf16_kernel
// (native/ops/sampling/sampling_kernels.cuh:55-200).
//
// Same one-wave streaming design as gemv_fp8.cu: 256-bit evict-first weight loads, four in
// flight per lane, activations staged per CTA in shared memory (bf16, bank-conflict-free
// planes), FHFMA.BF16 (fma.rn.f32.bf16: exact product, f32 accumulate) = 1 instruction per
// weight.  The argmax variant keeps a per-warp (value, index) best over the bf16-ROUNDED
// logits (what np.argmax sees after the reference's bf16 store), merges with a packed
// 64-bit atomicMax (value-order key | inverted index, so the lowest index wins ties) and the
// last CTA decodes the token and re-arms the scratch.
#include "common.cuh"
#include "gemv_common.cuh"
#include "../../include/pgk_b200.h"

#include <mutex>
#include <unordered_map>

namespace pgk {

template <int M>
__device__ __forceinline__ void stage_activations_bf16(const uint16_t* __restrict__ A, int K, uint16_t* a_s,
                                                       const uint16_t* __restrict__ gamma = nullptr, float eps = 0.f) {
    const int nvec = K >> 3, nchunk = K >> 4;
    float rinv[M];
#pragma unroll
    for (int m = 0; m < M; ++m)
        rinv[m] = gamma ? row_inv_rms_ref_order(A + (size_t)m * K, K, eps, reinterpret_cast<float*>(a_s)) : 1.f;
    const uint4* gv = reinterpret_cast<const uint4*>(gamma);
#pragma unroll
    for (int m = 0; m < M; ++m) {
        const uint4* src = reinterpret_cast<const uint4*>(A + (size_t)m * K);
        uint4* dst = reinterpret_cast<uint4*>(a_s + (size_t)m * K);
        for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
            uint4 v = ld_act16(src + j);
            if (gamma) v = norm8_bf16(v, __ldg(gv + j), rinv[m]);
            dst[(j & 1) * nchunk + (j >> 1)] = v;
        }
    }
    __syncthreads();
}

struct ArgmaxOut {
    unsigned long long* key;   // scratch, zero between launches
    unsigned* done;            // CTA arrival counter, zero between launches
    int32_t* token;            // result
};

template <int M, bool ARGMAX>
__global__ void __launch_bounds__(512)
gemv_bf16_kernel(const uint16_t* __restrict__ A, const uint16_t* __restrict__ B, uint16_t* __restrict__ C,
                 int K, int N, int rho, GemvEpilogue epi, ArgmaxOut am) {
    constexpr int U = 4;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint16_t* a_s = reinterpret_cast<uint16_t*>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    // rows dealt round-robin over the warps of the grid (see gemv_fp8.cu)
    const long gwarp = (long)blockIdx.x * wpc + warp;
    const long nwarps = (long)gridDim.x * wpc;
    const int row_begin = (int)min((long)N, gwarp);
    (void)rho;
    const int nchunk = K >> 4;

    u32x8 w[U];
    auto load_group = [&](int n, int cb) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            int c = cb + lane + 32 * u;
            if (c < nchunk) w[u] = ldg_stream256(B + (size_t)n * K + (size_t)c * 16);
        }
    };
    if (row_begin < N) load_group(row_begin, 0);
    griddep_wait();
    griddep_launch();
    stage_activations_bf16<M>(A, K, a_s, epi.norm_gamma, epi.norm_eps);

    const uint4* ap0 = reinterpret_cast<const uint4*>(a_s);
    bool preloaded = true;
    float best_v = -INFINITY;
    int best_i = 0x7fffffff;
    for (long nl = row_begin; nl < N; nl += nwarps) {
        const int n = (int)nl;
        float acc[M];
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] = 0.f;
        for (int cb = 0; cb < nchunk; cb += 32 * U) {
            if (!preloaded) load_group(n, cb);
            preloaded = false;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                int c = cb + lane + 32 * u;
                if (c < nchunk) {
#pragma unroll
                    for (int m = 0; m < M; ++m) {
                        const uint4* ap = ap0 + (size_t)m * (K >> 3);
                        float s0 = 0.f, s1 = 0.f;
                        dot8_bf16(s0, s1, &w[u].v[0], ap[c]);
                        dot8_bf16(s0, s1, &w[u].v[4], ap[nchunk + c]);
                        acc[m] += s0 + s1;
                    }
                }
            }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] = warp_sum(acc[m]);
        if (lane == 0) {
#pragma unroll
            for (int m = 0; m < M; ++m) {
                float v = acc[m];
                size_t o = (size_t)m * N + n;
                if (epi.bias) v = round_bf16(v) + bf16_bits_to_f32(epi.bias[n]);
                uint16_t ob = epi.residual ? hadd_bf16_bits(v, epi.residual[o]) : f32_to_bf16_bits(v);
                if (C) C[o] = ob;
                if (ARGMAX && m == 0) {
                    float r = bf16_bits_to_f32(ob);
                    if (r > best_v) { best_v = r; best_i = n; }   // rows ascend: first max kept
                }
            }
        }
    }
    if (ARGMAX) {
        __shared__ unsigned s_last;
        if (lane == 0 && best_i != 0x7fffffff) atomicMax(am.key, argmax_key(best_v, best_i));
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = (atomicAdd(am.done, 1u) == gridDim.x - 1);
        __syncthreads();
        if (s_last && threadIdx.x == 0) {
            __threadfence();
            unsigned long long k = atomicExch(am.key, 0ull);
            *am.token = 0x7fffffff - (int)(uint32_t)(k & 0xffffffffull);
            *am.done = 0;
        }
    }
}

template <bool ARGMAX>
__global__ void gemv_bf16_generic_kernel(const uint16_t* __restrict__ A, const uint16_t* __restrict__ B,
                                         uint16_t* __restrict__ C, int M, int K, int N, GemvEpilogue epi) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31;
    const long gw = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= (long)N * M) return;
    const int n = (int)(gw % N), m = (int)(gw / N);
    float acc = 0.f;
    for (int k = lane; k < K; k += 32)
        acc = fmaf(bf16_bits_to_f32(A[(size_t)m * K + k]), bf16_bits_to_f32(B[(size_t)n * K + k]), acc);
    acc = warp_sum(acc);
    if (lane == 0) {
        size_t o = (size_t)m * N + n;
        float v = acc;
        if (epi.bias) v = round_bf16(v) + bf16_bits_to_f32(epi.bias[n]);
        C[o] = epi.residual ? hadd_bf16_bits(v, epi.residual[o]) : f32_to_bf16_bits(v);
    }
}

static unsigned long long* g_am_scratch = nullptr;  // [0] key, [1] done counter

template <int M, bool ARGMAX>
static int launch_bf16(const uint16_t* A, const uint16_t* B, uint16_t* C, int K, int N, GemvEpilogue epi,
                       int32_t* token, cudaStream_t st) {
    size_t smem = (size_t)M * K * 2;
    auto kern = gemv_bf16_kernel<M, ARGMAX>;
    static std::mutex mu;
    static std::unordered_map<uint64_t, GemvCfg> cache;
    static bool attr_set = false;
    GemvCfg cfg;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!attr_set) {
            PGK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
            attr_set = true;
        }
        uint64_t key = ((uint64_t)(uint32_t)K << 32) | (uint32_t)N;
        auto it = cache.find(key);
        if (it == cache.end()) {
            int occ[17] = {0};
            for (int wpc = 2; wpc <= 16; ++wpc)
                PGK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[wpc], kern, wpc * 32, smem));
            it = cache.emplace(key, pick_gemv_cfg(N, occ)).first;
        }
        cfg = it->second;
    }
    ArgmaxOut am = {nullptr, nullptr, token};
    if (ARGMAX) {
        if (!g_am_scratch) {
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            cudaStreamIsCapturing(st, &cs);
            PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                        "gemv_bf16_argmax failed: first call must happen outside graph capture");
            PGK_CUDA(cudaMalloc(&g_am_scratch, 64));
            PGK_CUDA(cudaMemset(g_am_scratch, 0, 64));
        }
        am.key = g_am_scratch;
        am.done = reinterpret_cast<unsigned*>(g_am_scratch + 1);
    }
    PGK_CUDA(launch(kern, dim3(cfg.ctas), dim3(cfg.warps_per_cta * 32), smem, st, A, B, C, K, N, cfg.rho, epi, am));
    return 0;
}

static bool al(const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

static int bf16_dispatch(const void* A, const void* B, void* C, int M, int K, int N, const void* bias,
                         const void* residual, cudaStream_t st, const void* gamma = nullptr, float eps = 0.f) {
    PGK_REQUIRE(M >= 1 && K >= 1 && N >= 1, "gemv_bf16 failed: invalid shape M=%d K=%d N=%d", M, K, N);
    const uint16_t* a = (const uint16_t*)A;
    const uint16_t* b = (const uint16_t*)B;
    uint16_t* c = (uint16_t*)C;
    const bool fast = (K % 16 == 0) && al(A, 16) && al(B, 32) && ((size_t)K * 2 <= 200 * 1024);
    PGK_REQUIRE(!gamma || (fast && K >= 512 && al(gamma, 16)),
                "rmsnorm_gemv_bf16 failed: the fused norm needs K %% 16 == 0, K >= 512 and aligned pointers (K=%d)", K);
    if (!fast) {
        GemvEpilogue epi = {(const uint16_t*)bias, (const uint16_t*)residual};
        long warps = (long)N * M;
        PGK_CUDA(launch(gemv_bf16_generic_kernel<false>, dim3((unsigned)((warps * 32 + 255) / 256)), dim3(256),
                        0, st, a, b, c, M, K, N, epi));
        return 0;
    }
    int max_m = 4;
    while (max_m > 1 && (size_t)max_m * K * 2 > 200 * 1024) max_m >>= 1;
    for (int m0 = 0; m0 < M;) {
        int mt = M - m0 > max_m ? max_m : M - m0;
        if (mt == 3) mt = 2;
        GemvEpilogue epi = {(const uint16_t*)bias, residual ? (const uint16_t*)residual + (size_t)m0 * N : nullptr,
                            (const uint16_t*)gamma, eps};
        int rc;
        if (mt == 1)      rc = launch_bf16<1, false>(a + (size_t)m0 * K, b, c + (size_t)m0 * N, K, N, epi, nullptr, st);
        else if (mt == 2) rc = launch_bf16<2, false>(a + (size_t)m0 * K, b, c + (size_t)m0 * N, K, N, epi, nullptr, st);
        else              rc = launch_bf16<4, false>(a + (size_t)m0 * K, b, c + (size_t)m0 * N, K, N, epi, nullptr, st);
        if (rc) return rc;
        m0 += mt;
    }
    return 0;
}

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_gemv_bf16(const void* A, const void* B_nk, void* C, int M, int K, int N, void* stream) {
    return bf16_dispatch(A, B_nk, C, M, K, N, nullptr, nullptr, (cudaStream_t)stream);
}
int pgk_gemv_bf16_ex(const void* A, const void* B_nk, void* C, int M, int K, int N, const void* bias,
                     const void* residual, void* stream) {
    return bf16_dispatch(A, B_nk, C, M, K, N, bias, residual, (cudaStream_t)stream);
}
int pgk_gemv_bf16_argmax(const void* A, const void* B_nk, void* logits, int32_t* token, int K, int N,
                         void* stream) {
    PGK_REQUIRE(K % 16 == 0 && al(A, 16) && al(B_nk, 32) && (size_t)K * 2 <= 200 * 1024,
                "gemv_bf16_argmax failed: needs K%%16==0 and aligned pointers (K=%d)", K);
    GemvEpilogue epi = {nullptr, nullptr};
    return launch_bf16<1, true>((const uint16_t*)A, (const uint16_t*)B_nk, (uint16_t*)logits, K, N, epi,
                                token, (cudaStream_t)stream);
}

/* rmsnorm(x, gamma, eps) -> GEMV in one launch (norm rounded to bf16 exactly as the separate op). */
int pgk_rmsnorm_gemv_bf16(const void* A, const void* gamma, float eps, const void* B_nk, void* C, int M, int K,
                          int N, const void* bias, const void* residual, void* stream) {
    return bf16_dispatch(A, B_nk, C, M, K, N, bias, residual, (cudaStream_t)stream, gamma, eps);
}
int pgk_rmsnorm_gemv_bf16_argmax(const void* A, const void* gamma, float eps, const void* B_nk, void* logits,
                                 int32_t* token, int K, int N, void* stream) {
    PGK_REQUIRE(K % 16 == 0 && K >= 512 && al(A, 16) && al(B_nk, 32) && al(gamma, 16) && (size_t)K * 2 <= 200 * 1024,
                "rmsnorm_gemv_bf16_argmax failed: needs K%%16==0, K>=512 and aligned pointers (K=%d)", K);
    GemvEpilogue epi = {nullptr, nullptr, (const uint16_t*)gamma, eps};
    return launch_bf16<1, true>((const uint16_t*)A, (const uint16_t*)B_nk, (uint16_t*)logits, K, N, epi,
                                token, (cudaStream_t)stream);
}

}  // extern "C"
