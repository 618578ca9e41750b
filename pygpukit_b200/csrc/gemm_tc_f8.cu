// gemm_tc_f8.cu -- W8A16 skinny GEMM (decode batch 1..64) with the weights going to the tensor cores AS THEY ARE:
// raw E4M3 tiles by TMA (cp.async.bulk.tensor, 2-D tensor map, SWIZZLE_128B) -> tcgen05.mma.kind::f8f6f4 -> TMEM.
//
//   C[M,N] = A[M,K] . dequant(B[N,K])^T     B = FP8-E4M3 codes [N,K], S = bf16 [N/128, K/128], 1 <= M <= 64
//
// replaces gemv_fp8_warp_reduce_vec4_batched_kernel (fp8_opt.cuh:220-295) and the mma.sync path of
// pygpukit_w8a16_gemm_sm120 (w8a16_gemm.cu:104-339); on this repo's own path it supersedes gemm_tc.cu, whose converter
// warps turned every weight byte into BF16 on the CUDA cores (measured 1.2-1.9 TB/s of weight stream).
//
// Design: NO CUDA-core work on the weight stream.
//   * a pre-kernel (fp8_act_terms_kernel) scales every activation row by a power of two (max |a| -> [128, 256)) and
//     splits it into TWO E4M3 terms, a' = t0 + t1 / 16 with t0 = e4m3(a'), t1 = e4m3(16 (a' - t0)): BF16 has 8
//     significant bits, E4M3 has 4, so the two terms carry the value exactly down to 2^-10 of the row maximum
//     (smaller elements keep an absolute error <= 2^-18 of the maximum).  The terms are written once, as the
//     ready-to-use shared-memory image of the MMA's B operand (K-major rows of 128 bytes, 128-byte swizzle), one
//     [2 Mpad, 128 B] tile per 128-wide k block.
//   * main kernel, per CTA: 128 weight rows (MMA M = 128, one TMEM lane per row) x a k range (split-K fills the SMs).
//     Warp 0 = TMA producer: per k block one cp.async.bulk.tensor.2d (128 rows x 128 bytes of E4M3 codes, straight from
//     the [N, K] matrix, swizzled by the TMA unit) + one cp.async.bulk of the activation tile into a ring of stages,
//     completion on mbarriers.  Warp 1 = MMA issuer: 4 x tcgen05.mma.kind::f8f6f4 (128 x 2Mpad x 32) per k block from
//     shared-memory descriptors into one of two TMEM accumulator buffers, tcgen05.commit frees the stage and hands the
//     buffer over.  Warps 2-5 = promotion (DeepSeek-style two-level accumulation): tcgen05.ld the block's partial
//     products, acc[m] += scale[n/128, kb] * (D0[m] + D1[m] / 16) in FP32 registers -- this is where the 128x128 block
//     scale is applied and it bounds the tensor core's own accumulation to 128 products.
//   * epilogue: acc * 2^-shift[m], split-K partials folded in split order by the last-arriving CTA, bias / residual with
//     the reference's rounding points.
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/pgk_b200.h"

#include <cuda.h>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace pgk {

constexpr int kF8BK = 128;            // k per block = FP8 scale block = one 128-byte swizzle row
constexpr int kF8Threads = 192;       // TMA warp, MMA warp, 4 promotion warps
constexpr int kF8MaxKb = 256;

struct F8Params {
    const uint8_t* img;               // activation terms, [nkb][NB][128] swizzled
    const float* unshift;             // [Mpad] 2^-shift per activation row
    const uint16_t* S; uint16_t* C;
    const uint16_t* bias; const uint16_t* residual;
    int M, K, N, splits, mpad;
    float* ws; unsigned* counters;
};

// ---- activation rows -> two E4M3 terms, in the B operand's shared-memory image (one CTA per padded row)
__global__ void __launch_bounds__(256) fp8_act_terms_kernel(const uint16_t* __restrict__ A, uint8_t* __restrict__ img,
                                                            float* __restrict__ unshift, int M, int K, int mpad) {
    __shared__ uint32_t red[8];
    griddep_wait();
    griddep_launch();
    const int m = blockIdx.x, tid = threadIdx.x;
    const int NB = 2 * mpad, nkb = K / kF8BK;
    const uint16_t* row = A + (size_t)m * K;
    uint32_t mx = 0;
    if (m < M)
        for (int k = tid; k < K; k += blockDim.x) mx = max(mx, (uint32_t)(row[k] & 0x7fffu));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((tid & 31) == 0) red[tid >> 5] = mx;
    __syncthreads();
    mx = 0;
    for (int w = 0; w < 8; ++w) mx = max(mx, red[w]);
    int e = (int)(mx >> 7);                     // biased exponent of max |a|
    if (e == 255 || e == 0) e = 127;            // inf / nan / all-zero row: no shift
    e = min(max(e, 20), 234);
    const float up = __uint_as_float((uint32_t)(127 + 134 - e) << 23);      // 2^(134 - e): max |a| -> [128, 256)
    if (tid == 0) unshift[m] = (m < M) ? __uint_as_float((uint32_t)(127 - 134 + e) << 23) : 0.f;
    for (int k = tid; k < K; k += blockDim.x) {
        uint32_t c0 = 0, c1 = 0;
        if (m < M) {
            const float a = bf16_bits_to_f32(row[k]) * up;
            uint16_t p;
            asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(p) : "f"(0.f), "f"(a));
            c0 = p & 0xffu;
            const float r = (a - fp8_e4m3_to_f32(c0)) * 16.f;
            asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(p) : "f"(0.f), "f"(r));
            c1 = p & 0xffu;
        }
        const int kb = k >> 7, j = k & 127;
        uint8_t* tile = img + (size_t)kb * NB * 128;
        const int r0 = m, r1 = mpad + m;
        tile[r0 * 128 + ((((j >> 4) ^ (r0 & 7)) << 4) | (j & 15))] = (uint8_t)c0;
        tile[r1 * 128 + ((((j >> 4) ^ (r1 & 7)) << 4) | (j & 15))] = (uint8_t)c1;
    }
    (void)nkb;
}

// ---- descriptors
// shared-memory matrix descriptor, K-major, 128-byte swizzle: rows of 128 bytes, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: F32 accumulate, A = B = E4M3 (format 0), both K-major, N, M = 128
__host__ __device__ constexpr uint32_t umma_idesc_e4m3(int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);
}
__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint64_t pol) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;"
                 :: "r"(smem_dst), "l"(map), "r"(c0), "r"(c1), "r"(bar), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_load_1d(uint32_t smem_dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
// bounded: a wait that cannot complete (a pipeline bug, a lost TMA) traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    for (int spin = 0; !ok; ++spin) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && spin > (1 << 24)) __trap();
    }
}
__device__ __forceinline__ void umma_commit_u32(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}

// Several CTAs per SM (3 up to 16 batch rows, else 2): a CTA's fixed costs -- TMEM allocation, first TMA round trip, the
// split-K hand-over -- are as long as its stream of ~10 k blocks, so they must overlap another CTA's streaming.  Shared
// memory (4 / 3 stages) and TMEM (4 / 2 accumulator buffers: at most 512 columns per SM in total) are sized for that.
__host__ __device__ constexpr int f8_occ(int nb) { return nb <= 32 ? 3 : 2; }
__host__ __device__ constexpr int f8_stages(int nb) { return nb <= 64 ? 4 : 3; }
__host__ __device__ constexpr int f8_nbuf(int nb) { return nb <= 64 ? 4 : 2; }

template <int NB>      // NB = 2 * Mpad: MMA N (TMEM columns per accumulator buffer)
__global__ void __launch_bounds__(kF8Threads, f8_occ(NB))
gemm_fp8_f8mma_kernel(const __grid_constant__ CUtensorMap wmap, F8Params p) {
    constexpr int MP = NB / 2;
    constexpr uint32_t A_BYTES = kTcRows * kF8BK;               // 16 KB of E4M3 codes
    constexpr uint32_t B_BYTES = NB * 128;                      // activation terms
    constexpr uint32_t STAGE = A_BYTES + ((B_BYTES + 1023) & ~1023u);
    constexpr int S = f8_stages(NB);
    constexpr int NBUF = f8_nbuf(NB);
    constexpr uint32_t TMEM_COLS = NBUF * NB < 32 ? 32 : NBUF * NB;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t full_bar[8], empty_bar[8], dfull_bar[8], dempty_bar[8];
    __shared__ uint32_t tmem_base_s;
    __shared__ unsigned s_ticket;
    __shared__ float sc_s[kF8MaxKb];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, split = blockIdx.y;
    const int n0 = tile * kTcRows;
    const int nkb_total = p.K / kF8BK;
    const int per = (nkb_total + p.splits - 1) / p.splits;
    const int kb0 = split * per, kb1 = min(nkb_total, kb0 + per);
    const int nkb = max(0, kb1 - kb0);

    // block scales of this tile's k range: weights, read before the dependency wait
    for (int i = tid; i < nkb; i += kF8Threads) sc_s[i] = bf16_bits_to_f32(__ldg(p.S + (size_t)(n0 >> 7) * nkb_total + kb0 + i));
    if (warp == 1) tmem_alloc(&tmem_base_s, TMEM_COLS);
    if (tid == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int t = 0; t < NBUF; ++t) { mbar_init(&dfull_bar[t], 1); mbar_init(&dempty_bar[t], 4); }
        mbar_fence_init();
        asm volatile("prefetch.tensormap [%0];" :: "l"(&wmap) : "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_s;
    // the smem image must be 1024-byte aligned for the 128-byte swizzle
    const uint32_t smem_base = (smem_u32(smem) + 1023u) & ~1023u;

    if (warp == 0) {
        // ---------------- TMA producer
        if (lane == 0) {
            uint64_t pol;
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
            for (int i = 0; i < nkb; ++i) {
                const int s = i % S;
                const uint32_t fb = smem_u32(&full_bar[s]);
                mbar_wait_u32(smem_u32(&empty_bar[s]), (uint32_t)(((i / S) & 1) ^ 1));
                mbar_expect(fb, A_BYTES + B_BYTES);
                const uint32_t dst = smem_base + (uint32_t)s * STAGE;
                // weights never depend on the previous kernel; the activation image does (the pre-kernel wrote it)
                tma_load_2d(dst, &wmap, (kb0 + i) * kF8BK, n0, fb, pol);
                if (i == 0) griddep_wait();
                bulk_load_1d(dst + A_BYTES, p.img + (size_t)(kb0 + i) * B_BYTES, B_BYTES, fb);
            }
            if (nkb == 0) griddep_wait();
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_e4m3(NB);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % S, t = i % NBUF;
                mbar_wait_u32(smem_u32(&dempty_bar[t]), (uint32_t)(((i / NBUF) & 1) ^ 1));   // promotion drained this buffer
                mbar_wait_u32(smem_u32(&full_bar[s]), (uint32_t)((i / S) & 1));              // weights + terms landed
                tc_fence_after();
                const uint32_t a_addr = smem_base + (uint32_t)s * STAGE, b_addr = a_addr + A_BYTES;
                const uint64_t da = umma_desc_sw128(a_addr), db = umma_desc_sw128(b_addr);
#pragma unroll
                for (int ks = 0; ks < kF8BK / 32; ++ks)       // 32 bytes of k per MMA: +2 in the descriptor's 16-byte units
                    umma_f8(tmem_d + (uint32_t)t * NB, da + (uint64_t)(ks * 2), db + (uint64_t)(ks * 2), idesc, ks > 0 ? 1u : 0u);
                umma_commit_u32(smem_u32(&empty_bar[s]));     // stage free once these MMAs have read it
                umma_commit_u32(smem_u32(&dfull_bar[t]));     // block's partial products complete
            }
        }
    } else {
        // ---------------- promotion warps: TMEM lanes 32 * (warp % 4) .. + 31 = weight rows of the tile
        const int q = warp & 3;
        float acc[MP];
#pragma unroll
        for (int m = 0; m < MP; ++m) acc[m] = 0.f;
        for (int i = 0; i < nkb; ++i) {
            const int t = i % NBUF;
            mbar_wait_u32(smem_u32(&dfull_bar[t]), (uint32_t)((i / NBUF) & 1));
            tc_fence_after();
            const float sc = sc_s[i];
            const uint32_t base = tmem_d + ((uint32_t)(32 * q) << 16) + (uint32_t)t * NB;
            if (MP >= 16) {
#pragma unroll
                for (int c = 0; c < MP; c += 16) {
                    float d0[16], d1[16];
                    tmem_ld16(base + c, d0);
                    tmem_ld16(base + MP + c, d1);
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc[c + j] = fmaf(sc, fmaf(d1[j], 0.0625f, d0[j]), acc[c + j]);
                }
            } else {                                   // MP == 8: one load covers both terms
                float d[16];
                tmem_ld16(base, d);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fmaf(sc, fmaf(d[8 + j], 0.0625f, d[j]), acc[j]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&dempty_bar[t]);
        }
        griddep_wait();                                // bias / residual may come from the previous kernel
        // ---- epilogue
        const int row = 32 * q + lane;
        const int n = n0 + row;
        const size_t tile_ws = ((size_t)tile * p.splits + split) * (size_t)MP * kTcRows;
#pragma unroll
        for (int m = 0; m < MP; ++m) {
            if (m >= p.M) break;
            const float x0 = acc[m] * __ldg(p.unshift + m);
            if (p.splits == 1) {
                float x = x0;
                const size_t o = (size_t)m * p.N + n;
                if (p.bias) x = round_bf16(x) + bf16_bits_to_f32(p.bias[n]);
                p.C[o] = p.residual ? hadd_bf16_bits(x, p.residual[o]) : f32_to_bf16_bits(x);
            } else {
                __stcg(p.ws + tile_ws + (size_t)m * kTcRows + row, x0);
            }
        }
    }
    griddep_wait();                                    // every thread passes the wait before it releases the dependents
    griddep_launch();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_d, TMEM_COLS);
    if (p.splits == 1) return;
    if (tid == 0) {
        __threadfence();
        s_ticket = atomicAdd(p.counters + tile, 1u);
        __threadfence();
    }
    __syncthreads();
    if (s_ticket != (unsigned)(p.splits - 1)) return;
    // last CTA of the tile: fold the split-K partials in split order (deterministic), then the epilogue
    if (tid < kTcRows) {
        const int r = tid;
        const float* base = p.ws + (size_t)tile * p.splits * MP * kTcRows;
        const int nn = n0 + r;
        for (int m = 0; m < p.M; ++m) {
            float x = 0.f;
            for (int s0 = 0; s0 < p.splits; s0 += 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    v[u] = (s0 + u < p.splits) ? __ldcg(base + ((size_t)(s0 + u) * MP + m) * kTcRows + r) : 0.f;
#pragma unroll
                for (int u = 0; u < 8; ++u) x += v[u];
            }
            const size_t o = (size_t)m * p.N + nn;
            if (p.bias) x = round_bf16(x) + bf16_bits_to_f32(p.bias[nn]);
            p.C[o] = p.residual ? hadd_bf16_bits(x, p.residual[o]) : f32_to_bf16_bits(x);
        }
    }
    if (tid == 0) p.counters[tile] = 0;
}

// ---- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return (EncodeTiledFn)f;
    }();
    return fn;
}

static float* g_f8_ws = nullptr;
static unsigned* g_f8_cnt = nullptr;
static uint8_t* g_f8_img = nullptr;
static float* g_f8_unshift = nullptr;
static const size_t kF8WsBytes = 96ull << 20;
static const int kF8MaxTiles = 8192;
static const size_t kF8ImgBytes = (size_t)32768 * 128;      // K <= 32768, 2 * Mpad <= 128

template <int NB>
static int launch_f8(const CUtensorMap& map, F8Params p, cudaStream_t st) {
    constexpr uint32_t STAGE = kTcRows * kF8BK + ((NB * 128 + 1023) & ~1023u);
    constexpr int S = f8_stages(NB);
    constexpr size_t kSmem = (size_t)S * STAGE + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        PGK_CUDA(cudaFuncSetAttribute(gemm_fp8_f8mma_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        attr_set = true;
    }
    PGK_CUDA(launch(gemm_fp8_f8mma_kernel<NB>, dim3(p.N / kTcRows, p.splits), dim3(kF8Threads), kSmem, st, map, p));
    return 0;
}

bool gemm_fp8_f8_shape_ok(int M, int K, int N, const void* A, const void* B) {
    return get_encode() != nullptr && M >= 1 && M <= 64 && K % kF8BK == 0 && K / kF8BK <= kF8MaxKb && N % kTcRows == 0 &&
           N / kTcRows <= kF8MaxTiles && ((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0);
}

int gemm_fp8_f8(const void* A, const void* B_nk, const void* Sc, void* C, int M, int K, int N, const void* bias,
                const void* residual, cudaStream_t st) {
    PGK_REQUIRE(gemm_fp8_f8_shape_ok(M, K, N, A, B_nk),
                "gemm_fp8_f8 failed: needs 1 <= M <= 64, K %% 128 == 0, N %% 128 == 0, 16-byte aligned pointers and a driver "
                "with cuTensorMapEncodeTiled (M=%d K=%d N=%d)", M, K, N);
    if (!g_f8_ws) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(st, &cs);
        PGK_REQUIRE(cs == cudaStreamCaptureStatusNone, "gemm_fp8_f8 failed: first call must happen outside graph capture (workspace allocation)");
        PGK_CUDA(cudaMalloc(&g_f8_ws, kF8WsBytes));
        PGK_CUDA(cudaMalloc(&g_f8_cnt, kF8MaxTiles * sizeof(unsigned)));
        PGK_CUDA(cudaMemset(g_f8_cnt, 0, kF8MaxTiles * sizeof(unsigned)));
        PGK_CUDA(cudaMalloc(&g_f8_img, kF8ImgBytes));
        PGK_CUDA(cudaMalloc(&g_f8_unshift, 64 * sizeof(float)));
    }
    // tensor map of the weight matrix: [N rows, K bytes], box 128 x 128, 128-byte swizzle (cached per matrix)
    static std::mutex mu;
    static std::unordered_map<uint64_t, CUtensorMap> maps;
    CUtensorMap map;
    {
        std::lock_guard<std::mutex> lk(mu);
        const uint64_t key = (uint64_t)(uintptr_t)B_nk ^ ((uint64_t)K << 48) ^ ((uint64_t)N << 20);
        auto it = maps.find(key);
        if (it == maps.end()) {
            CUtensorMap m;
            const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)N};
            const cuuint64_t strides[1] = {(cuuint64_t)K};
            const cuuint32_t box[2] = {128, 128}, estr[2] = {1, 1};
            const CUresult r = get_encode()(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(B_nk), dims, strides, box, estr,
                                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            PGK_REQUIRE(r == CUDA_SUCCESS, "gemm_fp8_f8 failed: cuTensorMapEncodeTiled returned %d", (int)r);
            it = maps.emplace(key, m).first;
        }
        map = it->second;
    }
    const int mpad = M <= 8 ? 8 : (M <= 16 ? 16 : (M <= 32 ? 32 : 64));
    PGK_CUDA(launch(fp8_act_terms_kernel, dim3(mpad), dim3(256), 0, st, (const uint16_t*)A, g_f8_img, g_f8_unshift, M, K, mpad));
    const int tiles = N / kTcRows, nkb = K / kF8BK;
    // split-K so that one wave of resident CTAs (f8_occ per SM) covers the problem, every split keeping >= 2 k blocks
    const int slots = sm_count() * f8_occ(2 * mpad);
    int splits = slots / tiles;                            // floor: never a partial second wave
    if (splits > nkb / 2) splits = nkb / 2;
    if (splits > 16) splits = 16;
    if (splits < 1) splits = 1;
    while (splits > 1 && (size_t)tiles * splits * mpad * kTcRows * sizeof(float) > kF8WsBytes) --splits;
    F8Params p = {};
    p.img = g_f8_img; p.unshift = g_f8_unshift; p.S = (const uint16_t*)Sc; p.C = (uint16_t*)C;
    p.bias = (const uint16_t*)bias; p.residual = (const uint16_t*)residual;
    p.M = M; p.K = K; p.N = N; p.splits = splits; p.mpad = mpad; p.ws = g_f8_ws; p.counters = g_f8_cnt;
    if (mpad == 8) return launch_f8<16>(map, p, st);
    if (mpad == 16) return launch_f8<32>(map, p, st);
    if (mpad == 32) return launch_f8<64>(map, p, st);
    return launch_f8<128>(map, p, st);
}

}  // namespace pgk

extern "C" int pgk_gemm_fp8_bf16_f8mma(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N,
                                       const void* bias, const void* residual, void* stream) {
    bool clean = true;
    if (int rc = pgk::fp8_codes_clean(B_nk, (size_t)N * K, (cudaStream_t)stream, &clean)) return rc;
    PGK_REQUIRE(clean, "gemm_fp8_bf16_f8mma failed: the weights hold E4M3 codes 0x7F/0xFF (+-480 in the reference's tables, NaN for the "
                "tensor cores); use pgk_gemv_fp8_bf16, which serves them exactly");
    return pgk::gemm_fp8_f8(A, B_nk, S, C, M, K, N, bias, residual, (cudaStream_t)stream);
}
