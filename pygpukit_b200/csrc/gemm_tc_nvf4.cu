// gemm_tc_nvf4.cu -- NVF4 (W4A16) skinny GEMM for decode batch 5..64 on the 5th-gen tensor cores (tcgen05 + TMEM).
//
//   C[M,N] = alpha * A[M,K] . dequant(Bd[K/2,N], Bs[K/32,N])     Bd: two E2M1 nibbles per byte (low = even k),
//                                                                 Bs: UE4M3 scale per (32 k, column), 1 <= M <= 64
//
// replaces the batched use of gemv_nvf4_bf16_kernel (native/ops/matmul/gemv/w4a16_bf16/sm120/nvf4_kernels.cu:79-233: one
// launch per row, weights re-read M times) and the role of the CUTLASS NVF4 GEMM the reference cannot build here
// (native/ops/matmul/gemm/w4a16_bf16/sm120/nvf4_cutlass.cu).  On this repo's own path it replaces the slabs of four rows of
// gemv_nvf4.cu (weights re-read ceil(M/4) times) for M >= 5.
//
// Design: same weight-streaming pipeline as gemm_tc.cu (FP8).  The one difference is the dequantisation: an E2M1 value
// (<= 2 significant bits) times a UE4M3 scale (4 significant bits) has at most 6 significant bits and lies in
// [2^-8, 2880], so nibble x scale is EXACT in BF16 -- the converter warps write exact BF16 weight tiles, the
// activations go to the tensor cores as they are (BF16), and ONE tcgen05.mma.kind::f16 per 16 k computes exact
// products with F32 accumulation in TMEM: the same FP32-accumulate class as the CUDA-core GEMV, no hi/lo split and no
// block-scaled MMA (whose UE4M3 decoding would not reproduce the reference's "exponent 0 is normal" table,
// nvf4.cuh:65-115).
//   * a CTA owns 128 weight columns (MMA M = 128, one TMEM lane per column) and a k range (split-K over gridDim.y);
//     the batch is the MMA N (16 / 32 / 64 TMEM columns).
//   * per 128-wide k block: 64 byte rows x 128 B of nibbles + 4 x 128 B of scales stream global -> shared memory with
//     cp.async in full 128-byte lines (ring of raw blocks, requested before the dependency wait), thread (k chunk c,
//     column quad) gathers the four bytes of a column with PRMT, F2FP.F16.E2M1.UNPACK_B -> HMUL2 by the scale (exact
//     in F16) -> BF16x2, and stores the 16-byte k chunk into the canonical no-swizzle K-major core-matrix layout;
//     fence.proxy.async; a dedicated warp issues the 8 MMAs of the block and commits to the stage's mbarrier.
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

constexpr int kN4BK = 128;          // k per block (64 byte rows, 4 scale rows)
constexpr int kN4Threads = 512;     // 16 converter warps (+ one MMA-issuing warp)
constexpr int kN4Stages = 2;        // BF16 operand stages
constexpr int kN4RawW = 64 * 128 + 4 * 128;   // raw bytes of one weight block (nibbles + scales)
__host__ __device__ constexpr int n4_raw_depth(int nb) { return nb <= 16 ? 8 : (nb <= 32 ? 6 : 4); }

struct N4Params {
    const uint16_t* A; const uint8_t* Bd; const uint8_t* Bs; uint16_t* C;
    const uint16_t* bias; const uint16_t* residual;
    int M, K, N, splits;
    float alpha;
    float* ws; unsigned* counters;
};

template <int BYTE>
__device__ __forceinline__ uint32_t e2m1_pair_scaled_bf16x2(uint32_t q, __half2 s2) {
    const uint32_t h = e2m1x2_to_f16x2<BYTE>(q);
    const __half2 p = __hmul2(*reinterpret_cast<const __half2*>(&h), s2);      // exact: <= 6 significant bits
    const float2 f = __half22float2(p);
    return pack_bf16x2(f.x, f.y);                                               // exact: fits BF16
}

template <int NB>
__global__ void __launch_bounds__(kN4Threads + 32, 1)
gemm_nvf4_tc_kernel(N4Params p) {
    constexpr int NW = kN4Threads / 32;
    constexpr int RAW = n4_raw_depth(NB);
    constexpr int AI = (NB / 2 + NW - 1) / NW;             // activation items per thread per k block
    constexpr uint32_t A_STAGE = kTcRows * kN4BK * 2;      // 32 KB of BF16 weights
    constexpr uint32_t B_STAGE = NB * kN4BK * 2;           // BF16 activation tile
    constexpr uint32_t TMEM_COLS = NB < 32 ? 32 : NB;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t mma_bar[kN4Stages];
    __shared__ uint32_t tmem_base_s;
    __shared__ unsigned s_ticket;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, split = blockIdx.y;
    const int n0 = tile * kTcRows;
    const int nkb_total = p.K / kN4BK;
    const int per = (nkb_total + p.splits - 1) / p.splits;
    const int kb0 = split * per, kb1 = min(nkb_total, kb0 + per);
    const int nkb = max(0, kb1 - kb0);
    const int rr = lane & 7, piece = lane >> 3;
    const bool conv = warp < NW;

    if (warp == 0) tmem_alloc(&tmem_base_s, TMEM_COLS);
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kN4Stages; ++s) mbar_init(&mma_bar[s], 1);
        mbar_fence_init();
    }
    unsigned char* raw_w = smem + (size_t)kN4Stages * (A_STAGE + B_STAGE);
    unsigned char* raw_a = raw_w + (size_t)RAW * kN4RawW;
    // weights: 64 rows x 128 B of nibbles (one 16-byte piece per thread) + 4 rows x 128 B of scales (threads 0..31)
    auto issue_weights = [&](int kb, int rs) {
        unsigned char* dst0 = raw_w + (size_t)rs * kN4RawW;
        {
            const int j = tid >> 3, pc = tid & 7;
            const uint8_t* src = p.Bd + (size_t)(kb * 64 + j) * p.N + n0 + pc * 16;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(dst0 + j * 128 + pc * 16)), "l"(src) : "memory");
        }
        if (tid < 32) {
            const int j = tid >> 3, pc = tid & 7;
            const uint8_t* src = p.Bs + (size_t)(kb * 4 + j) * p.N + n0 + pc * 16;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(dst0 + 64 * 128 + j * 128 + pc * 16)), "l"(src) : "memory");
        }
    };
    auto issue_acts = [&](int kb, int rs) {
#pragma unroll
        for (int i = 0; i < AI; ++i) {
            const int m = (i * NW + warp) * 2 + (lane >> 4), pc = lane & 15;    // 2 rows x 256 B per warp request
            if (m < NB) {
                const bool ok = m < p.M;                       // rows past the batch are zero-filled (src-size 0)
                const uint16_t* src = p.A + (ok ? (size_t)m * p.K + (size_t)kb * kN4BK + pc * 8 : 0);
                const uint32_t dst = smem_u32(raw_a + (size_t)rs * (NB * kN4BK * 2) + m * 256 + ((pc ^ (m & 7)) * 16));
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
            }
        }
    };
    // group order as in gemm_tc.cu: W0..W(D-1) | A0..A(D-1) | (W_D, A_D) | ...
#pragma unroll
    for (int d = 0; d < RAW; ++d) {
        if (conv && d < nkb) issue_weights(kb0 + d, d);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    griddep_wait();
    griddep_launch();
#pragma unroll
    for (int d = 0; d < RAW; ++d) {
        if (conv && d < nkb) issue_acts(kb0 + d, d);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.wait_group %0;" :: "n"(RAW - 1) : "memory");       // block 0 (weights + slab) landed

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_s;
    constexpr uint32_t idesc = umma_idesc_bf16(NB);
    // converter role of this thread: k chunk c (8 k = 4 byte rows), columns 4 n4 .. 4 n4 + 3
    const int c = tid >> 5, n4 = tid & 31;

    for (int it = 0; it < nkb; ++it) {
        const int st = it & 1, kb = kb0 + it, rs = it % RAW;
        unsigned char* a_s = smem + (size_t)st * (A_STAGE + B_STAGE);
        unsigned char* b_s = a_s + A_STAGE;
        const unsigned char* rw = raw_w + (size_t)rs * kN4RawW;
        const unsigned char* ra = raw_a + (size_t)rs * (NB * kN4BK * 2);
        if (conv) {
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i] = *reinterpret_cast<const uint32_t*>(rw + (4 * c + i) * 128 + 4 * n4);
            const uint32_t sw = *reinterpret_cast<const uint32_t*>(rw + 64 * 128 + (c >> 2) * 128 + 4 * n4);
            uint4 aq[AI];
#pragma unroll
            for (int i = 0; i < AI; ++i) {
                const int q = i * NW + warp;                      // NB/2 items: row group q>>2, 64-byte quarter q&3
                if (q < NB / 2) {
                    const int m = 8 * (q >> 2) + rr, pc = (q & 3) * 4 + piece;
                    aq[i] = *reinterpret_cast<const uint4*>(ra + m * 256 + ((pc ^ rr) * 16));
                }
            }
            // the MMAs that read this BF16 stage two iterations ago must have completed
            if (it >= kN4Stages) mbar_wait(&mma_bar[st], (uint32_t)(((it >> 1) - 1) & 1));
            // weights: the four bytes of one column (k 8c .. 8c+7) x its block scale -> eight exact BF16 = one 16-byte k chunk
#pragma unroll
            for (int i2 = 0; i2 < 4; ++i2) {
                const int col = (i2 + (n4 >> 1)) & 3;             // rotated so that a quarter warp stores to 8 bank groups
                const uint32_t sel = (uint32_t)col | ((uint32_t)(4 + col) << 4);
                const uint32_t q = __byte_perm(__byte_perm(w[0], w[1], sel), __byte_perm(w[2], w[3], sel), 0x5410);
                const uint32_t sb = (sw >> (8 * col)) & 0x7fu;    // UE4M3 (bit 7 ignored, exponent 0 is normal) as F16 bits
                const uint32_t s16 = (sb << 7) + 0x2000u;
                const uint32_t s32 = s16 | (s16 << 16);
                const __half2 s2 = *reinterpret_cast<const __half2*>(&s32);
                uint4 o;
                o.x = e2m1_pair_scaled_bf16x2<0>(q, s2);
                o.y = e2m1_pair_scaled_bf16x2<1>(q, s2);
                o.z = e2m1_pair_scaled_bf16x2<2>(q, s2);
                o.w = e2m1_pair_scaled_bf16x2<3>(q, s2);
                const int n = 4 * n4 + col;
                *reinterpret_cast<uint4*>(a_s + (size_t)c * (kTcRows * 16) + (size_t)n * 16) = o;
            }
            // activations: plain copy of the k chunk into the canonical layout
#pragma unroll
            for (int i = 0; i < AI; ++i) {
                const int q = i * NW + warp;
                if (q < NB / 2) {
                    const int cc = (q & 3) * 4 + piece;
                    *reinterpret_cast<uint4*>(b_s + (size_t)cc * (NB * 16) + (size_t)(q >> 2) * 128 + rr * 16) = aq[i];
                }
            }
            fence_proxy_async();          // generic-proxy smem writes -> visible to the tensor core (async proxy)
            asm volatile("cp.async.wait_group %0;" :: "n"(RAW - 2) : "memory");   // this thread's part of block it+1 landed
        }
        __syncthreads();                  // BF16 stage complete; raw slot `rs` fully read; block it+1 visible to all
        if (conv) {
            if (it + RAW < nkb) { issue_weights(kb + RAW, rs); issue_acts(kb + RAW, rs); }
            asm volatile("cp.async.commit_group;" ::: "memory");
        } else if (lane == 0) {           // the MMA warp
            tc_fence_after();
            const uint32_t a_addr = smem_u32(a_s), b_addr = smem_u32(b_s);
#pragma unroll
            for (int ks = 0; ks < kN4BK / 16; ++ks) {
                const uint64_t da = umma_desc(a_addr + ks * 2 * (kTcRows * 16), kTcRows * 16, 128);
                const uint64_t db = umma_desc(b_addr + ks * 2 * (NB * 16), NB * 16, 128);
                umma_bf16(tmem_d, da, db, idesc, (it > 0 || ks > 0) ? 1u : 0u);
            }
            umma_commit(&mma_bar[st]);
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ---- epilogue
    if (nkb > 0) {
        const int last = nkb - 1;
        mbar_wait(&mma_bar[last & 1], (uint32_t)((last >> 1) & 1));
    }
    tc_fence_after();
    const int row = 32 * (warp & 3) + lane;                 // TMEM lane = weight column of the tile
    const int n = n0 + row;
    const size_t tile_ws = ((size_t)tile * p.splits + split) * (size_t)NB * kTcRows;
    if (warp < 4) {
#pragma unroll
        for (int cb = 0; cb < NB; cb += 16) {
            float v[16];
            if (nkb > 0) tmem_ld16(tmem_d + ((uint32_t)(32 * warp) << 16) + cb, v);
            else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int m = cb + j;
                if (m >= p.M) continue;
                if (p.splits == 1) {
                    float x = p.alpha * v[j];
                    const size_t o = (size_t)m * p.N + n;
                    if (p.bias) x = round_bf16(x) + bf16_bits_to_f32(p.bias[n]);
                    p.C[o] = p.residual ? hadd_bf16_bits(x, p.residual[o]) : f32_to_bf16_bits(x);
                } else {
                    __stcg(p.ws + tile_ws + (size_t)m * kTcRows + row, v[j]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_d, TMEM_COLS);
    if (p.splits == 1) return;
    if (tid == 0) {
        __threadfence();
        s_ticket = atomicAdd(p.counters + tile, 1u);
        __threadfence();
    }
    __syncthreads();
    if (s_ticket != (unsigned)(p.splits - 1)) return;
    // last CTA of the tile: fold the split-K partials in split order (deterministic), then the epilogue
    if (conv) {
        const int r = tid & (kTcRows - 1), half = tid >> 7;
        const float* base = p.ws + (size_t)tile * p.splits * NB * kTcRows;
        const int nn = n0 + r;
        for (int m = half; m < p.M; m += kN4Threads / kTcRows) {
            float x = 0.f;
            for (int s0 = 0; s0 < p.splits; s0 += 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    v[u] = (s0 + u < p.splits) ? __ldcg(base + ((size_t)(s0 + u) * NB + m) * kTcRows + r) : 0.f;
#pragma unroll
                for (int u = 0; u < 8; ++u) x += v[u];
            }
            x *= p.alpha;
            const size_t o = (size_t)m * p.N + nn;
            if (p.bias) x = round_bf16(x) + bf16_bits_to_f32(p.bias[nn]);
            p.C[o] = p.residual ? hadd_bf16_bits(x, p.residual[o]) : f32_to_bf16_bits(x);
        }
    }
    if (tid == 0) p.counters[tile] = 0;
}

static float* g_n4_ws = nullptr;
static unsigned* g_n4_cnt = nullptr;
static const size_t kN4WsBytes = 96ull << 20;
static const int kN4MaxTiles = 8192;

template <int NB>
static int launch_n4(N4Params p, cudaStream_t st) {
    const int tiles = p.N / kTcRows, nkb = p.K / kN4BK;
    const int sms = sm_count();
    int splits = 1;
    double best = -1.0;
    for (int s = 1; s <= 12 && s * 2 <= (nkb < 2 ? 2 : nkb); ++s) {
        const long ctas = (long)tiles * s;
        const long waves = (ctas + sms - 1) / sms;
        const int per = (nkb + s - 1) / s;
        const double t = (double)waves * (per + 2.0);      // whole waves; ~2 blocks of fixed cost per CTA
        const double score = 1.0 / t - 1e-6 * s;
        if (score > best) { best = score; splits = s; }
    }
    while (splits > 1 && (size_t)tiles * splits * NB * kTcRows * sizeof(float) > kN4WsBytes) --splits;
    if (!g_n4_ws) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(st, &cs);
        PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                    "gemm_nvf4_tc failed: first call must happen outside graph capture (workspace allocation)");
        PGK_CUDA(cudaMalloc(&g_n4_ws, kN4WsBytes));
        PGK_CUDA(cudaMalloc(&g_n4_cnt, kN4MaxTiles * sizeof(unsigned)));
        PGK_CUDA(cudaMemset(g_n4_cnt, 0, kN4MaxTiles * sizeof(unsigned)));
    }
    p.splits = splits;
    p.ws = g_n4_ws;
    p.counters = g_n4_cnt;
    constexpr size_t kSmem = (size_t)kN4Stages * (kTcRows * kN4BK * 2 + NB * kN4BK * 2) +
                             (size_t)n4_raw_depth(NB) * (kN4RawW + NB * kN4BK * 2);
    static bool attr_set = false;
    if (!attr_set) {
        PGK_CUDA(cudaFuncSetAttribute(gemm_nvf4_tc_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        attr_set = true;
    }
    PGK_CUDA(launch(gemm_nvf4_tc_kernel<NB>, dim3(tiles, splits), dim3(kN4Threads + 32), kSmem, st, p));
    return 0;
}

bool gemm_nvf4_tc_shape_ok(int M, int K, int N, const void* A, const void* Bd, const void* Bs) {
    return M >= 1 && M <= 64 && K % kN4BK == 0 && K >= kN4BK && N % kTcRows == 0 && N / kTcRows <= kN4MaxTiles &&
           ((uintptr_t)A % 16 == 0) && ((uintptr_t)Bd % 16 == 0) && ((uintptr_t)Bs % 16 == 0);
}

int gemm_nvf4_tc(const void* A, const void* Bd, const void* Bs, void* C, int M, int K, int N, float alpha, const void* bias,
                 const void* residual, cudaStream_t st) {
    PGK_REQUIRE(gemm_nvf4_tc_shape_ok(M, K, N, A, Bd, Bs),
                "gemm_nvf4_tc failed: needs 1 <= M <= 64, K %% 128 == 0, N %% 128 == 0 and 16-byte aligned pointers (M=%d K=%d N=%d)",
                M, K, N);
    N4Params p = {};
    p.A = (const uint16_t*)A; p.Bd = (const uint8_t*)Bd; p.Bs = (const uint8_t*)Bs; p.C = (uint16_t*)C;
    p.bias = (const uint16_t*)bias; p.residual = (const uint16_t*)residual;
    p.M = M; p.K = K; p.N = N; p.alpha = alpha;
    if (M <= 16) return launch_n4<16>(p, st);
    if (M <= 32) return launch_n4<32>(p, st);
    return launch_n4<64>(p, st);
}

}  // namespace pgk

extern "C" int pgk_gemm_nvf4_bf16_tc(const void* A, const void* B_data, const void* B_scale, void* C, int M, int K, int N,
                                     float alpha, const void* bias, const void* residual, void* stream) {
    return pgk::gemm_nvf4_tc(A, B_data, B_scale, C, M, K, N, alpha, bias, residual, (cudaStream_t)stream);
}
