// gemv_fp8.cu -- W8A16 GEMV for decode batch 1..8 on B200 (sm_100a), CUDA cores only.
//
//   C[M,N] = A[M,K] . dequant(B[N,K])^T      B = FP8-E4M3 codes, S = bf16 [N/128, K/128]
//
// replaces gemv_fp8_warp_reduce_vec4_kernel / _batched_kernel
// (native/ops/matmul/gemv/w8a16_bf16/sm120/fp8_opt.cuh:133-295).
//
// Design (HBM-bound; see DESIGN.md "gemv_fp8"):
//   * one wave: every warp of the grid is resident at once and owns `rho` consecutive
//     weight rows; the host picks (warps/CTA, CTAs/SM) so that N / (rho * warps) ~ 1.
//   * weights stream with 256-bit ld.global.nc.L1::no_allocate.L2::evict_first loads (LDG.E.NA.EFL2.256), four
//     256-bit loads in flight per lane; they are read ONCE for all M activation rows.
//   * dequantisation in registers: F2FP.F16.E4M3.UNPACK_B turns two codes into an exact
//     f16x2; the product with the activation is the mixed-precision FHFMA
//     (fma.rn.f32.f16: exact product, f32 accumulate) -- 1.5 instructions per weight.
//     Activations are staged once per CTA in shared memory as f16 after an exact
//     power-of-two rescale (bf16 has 8 significant bits, f16 holds 11; the per-row shift
//     puts max|a| at 2^14 so every element within 2^-28 of the maximum converts exactly).
//   * the 128x128 block scale is applied per 32-weight chunk in f32:
//     acc += scale * sum(chunk), so no rounding of LUT*scale is involved.
//   * epilogues keep the reference's rounding points (bias, residual add, SiLU*up).
//   * PDL: the first weight loads are issued before griddepcontrol.wait.
#include "common.cuh"
#include "gemv_common.cuh"
#include "../../include/pgk_b200.h"

#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace pgk {

// gemm_tc.cu: tcgen05 skinny GEMM for 16 <= M <= 64
bool gemm_fp8_tc_eligible(int M, int K, int N, const void* A, const void* B);
int gemm_fp8_tc(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N, const void* bias,
                const void* residual, cudaStream_t st);

// gemm_tc_f8.cu: raw E4M3 tiles by tensor-map TMA into tcgen05.mma.kind::f8f6f4 (PGK_TC_IMPL=2)
bool gemm_fp8_f8_shape_ok(int M, int K, int N, const void* A, const void* B);
int gemm_fp8_f8(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N, const void* bias,
                const void* residual, cudaStream_t st);
int gemm_fp8_tc_min_m();

constexpr int kUnroll = 4;

// ---- one-time scan of a weight tensor for the codes the hardware conversion decodes differently (see common.cuh)
__global__ void fp8_scan_codes_kernel(const uint8_t* __restrict__ B, size_t n, unsigned* __restrict__ flag) {
    const size_t nvec = n >> 4, tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    unsigned hit = 0;
    if ((reinterpret_cast<uintptr_t>(B) & 15) == 0) {
        for (size_t i = tid; i < nvec; i += nthr) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(B) + i);
            // a byte whose low seven bits are all ones carries into bit 7 when 1 is added (no carry crosses bytes)
            hit |= ((v.x & 0x7f7f7f7fu) + 0x01010101u) | ((v.y & 0x7f7f7f7fu) + 0x01010101u) |
                   ((v.z & 0x7f7f7f7fu) + 0x01010101u) | ((v.w & 0x7f7f7f7fu) + 0x01010101u);
        }
        for (size_t i = (nvec << 4) + tid; i < n; i += nthr) hit |= ((B[i] & 0x7fu) == 0x7fu) ? 0x80u : 0u;
    } else {
        for (size_t i = tid; i < n; i += nthr) hit |= ((B[i] & 0x7fu) == 0x7fu) ? 0x80u : 0u;
    }
    if (hit & 0x80808080u) atomicOr(flag, 1u);
}

static std::mutex g_scan_mu;
static std::unordered_map<const void*, bool> g_scan_clean;
static unsigned* g_scan_flag = nullptr;

int fp8_codes_clean(const void* codes, size_t nbytes, cudaStream_t st, bool* clean) {
    std::lock_guard<std::mutex> lk(g_scan_mu);
    auto it = g_scan_clean.find(codes);
    if (it != g_scan_clean.end()) { *clean = it->second; return 0; }
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cs);
    PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                "gemv_fp8_bf16 failed: the first use of an FP8 weight tensor must happen outside graph capture (one-time code scan)");
    if (!g_scan_flag) PGK_CUDA(cudaMalloc(&g_scan_flag, sizeof(unsigned)));
    PGK_CUDA(cudaMemsetAsync(g_scan_flag, 0, sizeof(unsigned), st));
    const size_t want = (nbytes / 16 + 255) / 256;
    const unsigned grid = (unsigned)(want < 1 ? 1 : (want > 4096 ? 4096 : want));
    fp8_scan_codes_kernel<<<grid, 256, 0, st>>>((const uint8_t*)codes, nbytes, g_scan_flag);
    PGK_LAUNCH_CHECK("fp8_scan_codes");
    unsigned h = 0;
    PGK_CUDA(cudaMemcpyAsync(&h, g_scan_flag, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    PGK_CUDA(cudaStreamSynchronize(st));
    *clean = (h == 0);
    g_scan_clean.emplace(codes, *clean);
    return 0;
}
void fp8_codes_forget(const void* base) {
    std::lock_guard<std::mutex> lk(g_scan_mu);
    g_scan_clean.erase(base);
}

// NMAT = 1: plain GEMV; NMAT = 2: gate/up pair with SiLU*up epilogue.
//
// One group (U x 256-bit loads per lane) in flight per warp, up to 16 warps per CTA.  (A register double-buffered
// form was measured slower on B200 -- 168 registers cap the CTA at 12 warps -- see profiles/r01_*; removed.)
// XCHG: the tensor-parallel form with the all-reduce inside (a separate instantiation: the plain kernel pays nothing for it)
template <int M, int NMAT, bool XCHG = false>
__global__ void __launch_bounds__(512)
gemv_fp8_sb_kernel(const uint16_t* __restrict__ A, const uint8_t* __restrict__ B0,
                const uint16_t* __restrict__ S0, const uint8_t* __restrict__ B1,
                const uint16_t* __restrict__ S1, uint16_t* __restrict__ C, int K, int N, int rho,
                GemvEpilogue epi) {
    constexpr int U = (NMAT == 2) ? 2 : kUnroll;   // 256-bit loads in flight per lane per matrix
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __half* a_s = reinterpret_cast<__half*>(smem_raw);
    float* inv_s = reinterpret_cast<float*>(smem_raw + (size_t)M * K * 2);
    uint32_t* red = reinterpret_cast<uint32_t*>(inv_s + 8);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpc = blockDim.x >> 5;
    const long gwarp = (long)blockIdx.x * wpc + warp;
    const long nwarps = (long)gridDim.x * wpc;
    const int row_begin = (int)min((long)N, gwarp);
    (void)rho;
    const int nchunk = K >> 5;
    const int SK = (K + 127) >> 7;

    u32x8 w[NMAT][U];
    uint16_t wsc[NMAT][U];     // the 128x128 block scales ride with their weights (same latency window)
    auto load_group = [&](int n, int cb) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            int c = cb + lane + 32 * u;
            if (c < nchunk) {
                w[0][u] = ldg_stream256(B0 + (size_t)n * K + (size_t)c * 32);
                wsc[0][u] = __ldg(S0 + (size_t)(n >> 7) * SK + (c >> 2));
                if (NMAT == 2) {
                    w[NMAT - 1][u] = ldg_stream256(B1 + (size_t)n * K + (size_t)c * 32);
                    wsc[NMAT - 1][u] = __ldg(S1 + (size_t)(n >> 7) * SK + (c >> 2));
                }
            }
        }
    };
    // weights never depend on the previous kernel: start pulling the first row now
    if (row_begin < N) load_group(row_begin, 0);
    griddep_wait();
    griddep_launch();
    const bool xchg = XCHG && M == 1 && NMAT == 1 && epi.xchg.world > 1;
    const unsigned xseq = xchg ? *epi.xchg.epoch + 1u : 0u;
    unsigned xticket = 0;
    if (xchg && threadIdx.x == 0) xticket = xchg_arrive(epi.xchg);
    __shared__ uint16_t xrow[XCHG ? 1024 : 1];              // exchange: the CTA's partials, [row iteration][warp]
    __shared__ int xcnt[XCHG ? 64 : 1];                     // exchange: warps of the CTA done with row iteration i (N <= 8192: <= 28 iterations)
    int xit = 0;
    if (xchg && threadIdx.x < 64) xcnt[threadIdx.x] = 0;       // visible after the barriers of the staging below
    stage_activations_f16<M>(A, K, a_s, inv_s, red, epi.norm_gamma, epi.norm_eps);

    const uint4* a_lo = reinterpret_cast<const uint4*>(a_s);
    bool preloaded = true;
    for (long nl = row_begin; nl < N; nl += nwarps) {
        const int n = (int)nl;
        float acc[NMAT][M];
#pragma unroll
        for (int t = 0; t < NMAT; ++t)
#pragma unroll
            for (int m = 0; m < M; ++m) acc[t][m] = 0.f;
        // the residual of this row is requested now and consumed after the dot product
        uint16_t res[M];
        if (NMAT == 1 && epi.residual && lane == 0) {
#pragma unroll
            for (int m = 0; m < M; ++m) res[m] = epi.residual[(size_t)m * N + n];
        }
        for (int cb = 0; cb < nchunk; cb += 32 * U) {
            if (!preloaded) load_group(n, cb);
            preloaded = false;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                int c = cb + lane + 32 * u;
                if (c < nchunk) {
                    float sc0 = bf16_bits_to_f32(wsc[0][u]);
                    float sc1 = (NMAT == 2) ? bf16_bits_to_f32(wsc[NMAT - 1][u]) : 0.f;
#pragma unroll
                    for (int m = 0; m < M; ++m) {
                        const uint4* ap = a_lo + (size_t)m * (K >> 3);
                        acc[0][m] = fmaf(dot32_e4m3(w[0][u], ap, nchunk, c), sc0, acc[0][m]);
                        if (NMAT == 2)
                            acc[NMAT - 1][m] = fmaf(dot32_e4m3(w[NMAT - 1][u], ap, nchunk, c), sc1, acc[NMAT - 1][m]);
                    }
                }
            }
        }
#pragma unroll
        for (int t = 0; t < NMAT; ++t)
#pragma unroll
            for (int m = 0; m < M; ++m) acc[t][m] = warp_sum(acc[t][m]);
        if (lane == 0) {
#pragma unroll
            for (int m = 0; m < M; ++m) {
                float v = acc[0][m] * inv_s[m];
                size_t o = (size_t)m * N + n;
                if (NMAT == 2) {
                    float g = round_bf16(v);
                    float u = round_bf16(acc[NMAT - 1][m] * inv_s[m]);
                    float s = round_bf16(silu_ref(g));
                    C[o] = f32_to_bf16_bits(s * u);
                } else if (xchg) {
                    xrow[xit * wpc + warp] = f32_to_bf16_bits(v);
                } else {
                    if (epi.bias) v = round_bf16(v) + bf16_bits_to_f32(epi.bias[n]);
                    C[o] = epi.residual ? hadd_bf16_bits(v, res[m]) : f32_to_bf16_bits(v);
                }
            }
        }
        if (xchg) {
            // Row-parallel layer under tensor parallelism.  In row iteration i the warps of this CTA own the consecutive
            // rows i * nwarps + blockIdx.x * wpc + (0 .. wpc-1).  The warp that completes the group publishes it to every
            // rank as contiguous LL words (one coalesced store per rank instead of wpc scattered 8-byte NVLink writes)
            // while the other rows are still being computed: the NVLink hop hides behind the rest of the GEMV.
            const int base = (int)((long)xit * nwarps + (long)blockIdx.x * wpc);
            const int expect = min(wpc, N - base);
            int last = 0;
            if (lane == 0) {
                __threadfence_block();
                last = atomicAdd(&xcnt[xit], 1) == expect - 1;
            }
            last = __shfl_sync(0xffffffffu, last, 0);
            if (last) {
                __threadfence_block();
                if (lane < expect) {
                    const uint16_t mine = *reinterpret_cast<volatile uint16_t*>(&xrow[xit * wpc + lane]);
                    for (int r = 0; r < epi.xchg.world; ++r)
                        if (r != epi.xchg.rank) xchg_publish(epi.xchg, xseq, base + lane, mine, r);
                }
            }
        }
        ++xit;
    }
    if (xchg) {
        // every group of this CTA has been published: sum the ranks' words for the same rows in rank order (+ residual)
        // -- pgk_p2p_ll_allreduce_add_bf16's arithmetic
        __syncthreads();
        const int iters = (int)((N - (long)blockIdx.x * wpc + nwarps - 1) / nwarps);       // row iterations of this CTA (<= 0: none)
        for (int i = warp; i < iters; i += wpc) {
            const int n = (int)((long)i * nwarps + (long)blockIdx.x * wpc + lane);
            if (lane < wpc && n < N) {
                float acc = 0.f;
                for (int r = 0; r < epi.xchg.world; ++r)
                    acc += (r == epi.xchg.rank) ? bf16_bits_to_f32(xrow[i * wpc + lane]) : xchg_fetch(epi.xchg, xseq, n, r);
                C[n] = epi.residual ? hadd_bf16_bits(acc, epi.residual[n]) : f32_to_bf16_bits(acc);
            }
        }
        if (threadIdx.x == 0) xchg_finish(epi.xchg, xseq, xticket, gridDim.x);
    }
}


// Generic fallback (any K, any alignment): warp per row, byte loads, exact LUT semantics
// including 0x7F/0xFF -> +-480 (fp8.cuh:89,121).
template <int NMAT>
__global__ void gemv_fp8_generic_kernel(const uint16_t* __restrict__ A, const uint8_t* __restrict__ B0,
                                        const uint16_t* __restrict__ S0, const uint8_t* __restrict__ B1,
                                        const uint16_t* __restrict__ S1, uint16_t* __restrict__ C,
                                        int M, int K, int N, GemvEpilogue epi) {
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31;
    const long gw = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= (long)N * M) return;
    const int n = (int)(gw % N), m = (int)(gw / N);
    const int SK = (K + 127) >> 7;
    float acc0 = 0.f, acc1 = 0.f;
    for (int k = lane; k < K; k += 32) {
        float a = bf16_bits_to_f32(A[(size_t)m * K + k]);
        float s = bf16_bits_to_f32(S0[(size_t)(n >> 7) * SK + (k >> 7)]);
        acc0 = fmaf(a, fp8_e4m3_to_f32(B0[(size_t)n * K + k]) * s, acc0);
        if (NMAT == 2) {
            float s2 = bf16_bits_to_f32(S1[(size_t)(n >> 7) * SK + (k >> 7)]);
            acc1 = fmaf(a, fp8_e4m3_to_f32(B1[(size_t)n * K + k]) * s2, acc1);
        }
    }
    acc0 = warp_sum(acc0);
    if (NMAT == 2) acc1 = warp_sum(acc1);
    if (lane == 0) {
        size_t o = (size_t)m * N + n;
        float v = acc0;
        if (NMAT == 2) {
            float g = round_bf16(v), u = round_bf16(acc1);
            C[o] = f32_to_bf16_bits(round_bf16(silu_ref(g)) * u);
        } else {
            if (epi.bias) v = round_bf16(v) + bf16_bits_to_f32(epi.bias[n]);
            C[o] = epi.residual ? hadd_bf16_bits(v, epi.residual[o]) : f32_to_bf16_bits(v);
        }
    }
}

template <typename Kern>
static int launch_fp8_kern(Kern kern, int maxw, int M, const void* A, const void* B0, const void* S0, const void* B1,
                           const void* S1, void* C, int K, int N, GemvEpilogue epi, cudaStream_t st,
                           std::unordered_map<uint64_t, GemvCfg>& cache, bool& attr_set) {
    size_t smem = (size_t)M * K * 2 + 8 * sizeof(float) + 8 * 16 * sizeof(uint32_t);
    static std::mutex mu;
    GemvCfg cfg;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!attr_set) {
            PGK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 222 * 1024));     // + 2.3 KB static (xrow, xcnt) + the 1 KB the system reserves <= 227 KB
            attr_set = true;
        }
        uint64_t key = ((uint64_t)(uint32_t)K << 32) | (uint32_t)N;
        auto it = cache.find(key);
        if (it == cache.end()) {
            int occ[17] = {0};
            for (int wpc = 2; wpc <= maxw; ++wpc) {
                int nb = 0;
                PGK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, wpc * 32, smem));
                occ[wpc] = nb;
            }
            PGK_REQUIRE(occ[4] > 0, "gemv_fp8_bf16 failed: kernel does not fit (K=%d, M=%d)", K, M);
            it = cache.emplace(key, pick_gemv_cfg(N, occ)).first;
        }
        cfg = it->second;
    }
    PGK_CUDA(launch(kern, dim3(cfg.ctas), dim3(cfg.warps_per_cta * 32), smem, st,
                    (const uint16_t*)A, (const uint8_t*)B0, (const uint16_t*)S0, (const uint8_t*)B1,
                    (const uint16_t*)S1, (uint16_t*)C, K, N, cfg.rho, epi));
    return 0;
}

template <int M, int NMAT, bool XCHG = false>
static int launch_fp8(const void* A, const void* B0, const void* S0, const void* B1, const void* S1,
                      void* C, int K, int N, GemvEpilogue epi, cudaStream_t st) {
    static std::unordered_map<uint64_t, GemvCfg> cache;
    static bool attr = false;
    return launch_fp8_kern(gemv_fp8_sb_kernel<M, NMAT, XCHG>, 16, M, A, B0, S0, B1, S1, C, K, N, epi, st, cache, attr);
}

// with peers open every M = 1 GEMV runs the XCHG instantiation (exchange off: world = 0), so that a tensor-parallel step
// keeps ONE hot GEMV kernel in the instruction caches (measured: gemv_nvf4.cu::tp_active)
static bool tp_active_fp8() {
    GemvXchg x;
    return p2p_gemv_xchg(&x) && x.world > 1;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; }

// Dispatch over M (activation rows).  M > 4 is processed in slabs of <= 4 rows (weights are
// re-read once per slab; the tcgen05 skinny GEMM takes over for M >= 16, see gemm_w8a16_tc.cu).
static int gemv_fp8_dispatch(const void* A, const void* B0, const void* S0, const void* B1,
                             const void* S1, void* C, int M, int K, int N, const void* bias,
                             const void* residual, bool swiglu, cudaStream_t st, const void* gamma = nullptr,
                             float eps = 0.f) {
    PGK_REQUIRE(M >= 1 && K >= 1 && N >= 1, "gemv_fp8_bf16 failed: invalid shape M=%d K=%d N=%d", M, K, N);
    bool fast = (K % 32 == 0) && aligned16(A) && aligned32(B0) && (!swiglu || aligned32(B1)) &&
                ((size_t)K * 2 + 1024 <= 200 * 1024);
    // weights holding codes 0x7F / 0xFF keep the reference's table semantics (+-480): exact kernel, never the cvt path
    bool clean = true;
    if (int rc = fp8_codes_clean(B0, (size_t)N * K, st, &clean)) return rc;
    if (clean && swiglu) { if (int rc = fp8_codes_clean(B1, (size_t)N * K, st, &clean)) return rc; }
    PGK_REQUIRE(clean || !gamma, "rmsnorm_gemv_fp8_bf16 failed: weights with E4M3 codes 0x7F/0xFF take the exact path, "
                "which has no fused norm (run pgk_rmsnorm first)");
    if (!clean) fast = false;
    const uint16_t* a = (const uint16_t*)A;
    uint16_t* c = (uint16_t*)C;
    PGK_REQUIRE(!gamma || (fast && K >= 512 && aligned16(gamma)),
                "rmsnorm_gemv_fp8_bf16 failed: the fused norm needs K %% 32 == 0, K >= 512 and aligned pointers (K=%d)", K);
    // decode batch >= 5 (PGK_TC_MIN_M): the weight-streaming tensor-core GEMM (PGK_TC=0 keeps the slab GEMV for A/B timing)
    static const bool tc_on = !(getenv("PGK_TC") && atoi(getenv("PGK_TC")) == 0);
    static const int tc_impl = getenv("PGK_TC_IMPL") ? atoi(getenv("PGK_TC_IMPL")) : 1;
    if (tc_on && clean && !swiglu && !gamma && tc_impl == 2 && M >= gemm_fp8_tc_min_m() && gemm_fp8_f8_shape_ok(M, K, N, A, B0))
        return gemm_fp8_f8(A, B0, S0, C, M, K, N, bias, residual, st);
    if (tc_on && clean && !swiglu && !gamma && gemm_fp8_tc_eligible(M, K, N, A, B0))
        return gemm_fp8_tc(A, B0, S0, C, M, K, N, bias, residual, st);
    if (!fast) {
        GemvEpilogue epi = {(const uint16_t*)bias, (const uint16_t*)residual};
        long warps = (long)N * M;
        int block = 256;
        long grid = (warps * 32 + block - 1) / block;
        if (swiglu)
            PGK_CUDA(launch(gemv_fp8_generic_kernel<2>, dim3((unsigned)grid), dim3(block), 0, st, a,
                            (const uint8_t*)B0, (const uint16_t*)S0, (const uint8_t*)B1,
                            (const uint16_t*)S1, c, M, K, N, epi));
        else
            PGK_CUDA(launch(gemv_fp8_generic_kernel<1>, dim3((unsigned)grid), dim3(block), 0, st, a,
                            (const uint8_t*)B0, (const uint16_t*)S0, (const uint8_t*)nullptr,
                            (const uint16_t*)nullptr, c, M, K, N, epi));
        return 0;
    }
    // rows per slab limited by shared memory (M*K*2 bytes) and registers
    int max_m = 4;
    while (max_m > 1 && (size_t)max_m * K * 2 + 1024 > 200 * 1024) max_m >>= 1;
    if (swiglu && max_m > 2) max_m = 2;
    for (int m0 = 0; m0 < M;) {
        int mt = M - m0;
        if (mt > max_m) mt = max_m;
        GemvEpilogue epi = {(const uint16_t*)bias,
                            residual ? (const uint16_t*)residual + (size_t)m0 * N : nullptr,
                            (const uint16_t*)gamma, eps};
        const uint16_t* am = a + (size_t)m0 * K;
        uint16_t* cm = c + (size_t)m0 * N;
        int rc = 0;
        if (swiglu) {
            if (mt == 1) rc = launch_fp8<1, 2>(am, B0, S0, B1, S1, cm, K, N, epi, st);
            else         rc = launch_fp8<2, 2>(am, B0, S0, B1, S1, cm, K, N, epi, st);
        } else {
            if (mt == 1 && tp_active_fp8()) rc = launch_fp8<1, 1, true>(am, B0, S0, nullptr, nullptr, cm, K, N, epi, st);
            else if (mt == 1) rc = launch_fp8<1, 1>(am, B0, S0, nullptr, nullptr, cm, K, N, epi, st);
            else if (mt == 2) rc = launch_fp8<2, 1>(am, B0, S0, nullptr, nullptr, cm, K, N, epi, st);
            else if (mt == 3) rc = launch_fp8<3, 1>(am, B0, S0, nullptr, nullptr, cm, K, N, epi, st);
            else              rc = launch_fp8<4, 1>(am, B0, S0, nullptr, nullptr, cm, K, N, epi, st);
        }
        if (rc) return rc;
        m0 += mt;
    }
    return 0;
}

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_gemv_fp8_bf16(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N,
                      void* stream) {
    return gemv_fp8_dispatch(A, B_nk, S, nullptr, nullptr, C, M, K, N, nullptr, nullptr, false,
                             (cudaStream_t)stream);
}

int pgk_gemv_fp8_bf16_ex(const void* A, const void* B_nk, const void* S, void* C, int M, int K,
                         int N, const void* bias, const void* residual, void* stream) {
    return gemv_fp8_dispatch(A, B_nk, S, nullptr, nullptr, C, M, K, N, bias, residual, false,
                             (cudaStream_t)stream);
}

int pgk_gemv_fp8_swiglu_bf16(const void* A, const void* Wg_nk, const void* Sg, const void* Wu_nk,
                             const void* Su, void* C, int M, int K, int N, void* stream) {
    return gemv_fp8_dispatch(A, Wg_nk, Sg, Wu_nk, Su, C, M, K, N, nullptr, nullptr, true,
                             (cudaStream_t)stream);
}

/* Row-parallel linear of ONE token under tensor parallelism: out = bf16(sum over ranks of bf16(x_r . W_r^T)) + residual in ONE
 * launch (out may alias residual) -- pgk_gemv_fp8_bf16 followed by pgk_p2p_ll_allreduce_add_bf16, bit for bit, without
 * the second launch.  Needs the peer buffers (pgk_p2p_open), N <= 8192 and the streaming kernel's shape rules. */
int pgk_gemv_fp8_allreduce_bf16(const void* A, const void* B_nk, const void* S, const void* residual, void* out, int K,
                                int N, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    GemvEpilogue epi = {nullptr, (const uint16_t*)residual};
    PGK_REQUIRE(p2p_gemv_xchg(&epi.xchg), "gemv_fp8_allreduce failed: peer buffers not opened (pgk_p2p_create / pgk_p2p_open)");
    bool clean = true;
    if (int rc = fp8_codes_clean(B_nk, (size_t)N * K, st, &clean)) return rc;
    PGK_REQUIRE(clean && K % 32 == 0 && aligned16(A) && aligned32(B_nk) && (size_t)K * 2 + 1024 <= 200 * 1024 && N >= 1 &&
                (unsigned)N <= epi.xchg.slot_words,
                "gemv_fp8_allreduce failed: needs K %% 32 == 0, N <= %u, aligned pointers and weights without the codes 0x7F / 0xFF "
                "(K=%d N=%d)", epi.xchg.slot_words, K, N);
    if (epi.xchg.world <= 1) epi.xchg.world = 0;
    PGK_REQUIRE(N <= 8192, "gemv_fp8_allreduce failed: N=%d", N);      // xrow / xcnt: <= 28 row iterations of <= 16 warps per CTA
    return launch_fp8<1, 1, true>((const uint16_t*)A, B_nk, S, nullptr, nullptr, (uint16_t*)out, K, N, epi, st);
}

/* rmsnorm(x, gamma, eps) -> W8A16 GEMV in ONE launch: every CTA normalises the activation rows while it
 * stages them (same summation order and bf16 rounding as pgk_rmsnorm / norm_kernels.cuh:535-585). */
int pgk_rmsnorm_gemv_fp8_bf16(const void* A, const void* gamma, float eps, const void* B_nk, const void* S,
                              void* C, int M, int K, int N, const void* bias, const void* residual, void* stream) {
    return gemv_fp8_dispatch(A, B_nk, S, nullptr, nullptr, C, M, K, N, bias, residual, false,
                             (cudaStream_t)stream, gamma, eps);
}
int pgk_rmsnorm_gemv_fp8_swiglu_bf16(const void* A, const void* gamma, float eps, const void* Wg_nk,
                                     const void* Sg, const void* Wu_nk, const void* Su, void* C, int M, int K,
                                     int N, void* stream) {
    return gemv_fp8_dispatch(A, Wg_nk, Sg, Wu_nk, Su, C, M, K, N, nullptr, nullptr, true,
                             (cudaStream_t)stream, gamma, eps);
}

}  // extern "C"
