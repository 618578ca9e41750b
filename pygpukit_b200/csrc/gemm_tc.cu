// gemm_tc.cu -- W8A16 skinny GEMM for decode batch 16..64 on the 5th-gen tensor cores (tcgen05 + TMEM).
//
//   C[M,N] = A[M,K] . dequant(B[N,K])^T     B = FP8-E4M3 codes [N,K], S = bf16 [N/128, K/128], 1 <= M <= 64 (used for M >= 5)
//
// replaces gemv_fp8_warp_reduce_vec4_batched_kernel (fp8_opt.cuh:220-295, re-reads the weights M times) and
// the M >= 16 tensor-core path of pygpukit_w8a16_gemm_sm120 (w8a16_gemm.cu:104-339: dequantise to BF16 in
// shared memory, mma.sync m16n8k16, unpipelined).
//
// Design (DESIGN.md "gemm_fp8_tc"): the problem is still weight-streaming bound (M <= 64 rows share one read
// of N*K bytes), so the kernel is organised around the HBM stream and hands the arithmetic to tcgen05:
//   * the GEMM is computed transposed, D^T[n, m] = W[n, :] . A[m, :]: a CTA owns 128 weight rows (the MMA's
//     M = 128 -> one TMEM lane per weight row) and the batch is the MMA's N (16 / 32 / 64 f32 TMEM columns);
//     split-K over gridDim.y fills the 148 SMs when N/128 tiles are too few.
//   * per 128-wide k block (= one FP8 scale block; the 128-row tile lies inside one scale row, so the CTA sees
//     ONE scale per k block): 16 KB of E4M3 codes stream global -> shared memory with cp.async (LDGSTS, a ring of
//     four raw blocks = 64 KB in flight per SM, requested before the dependency wait), are converted exactly to BF16
//     (F2FP.F16.E4M3.UNPACK_B -> f32 -> BF16x2 pack; E4M3 fits BF16) and stored to shared memory in the
//     canonical no-swizzle K-major core-matrix layout; the [M,128] activation slab is multiplied by the block
//     scale in f32 and split into TWO BF16 terms, hi = bf16(a*s) and lo = bf16(a*s - hi), stored in the same
//     layout: with exact BF16 weights the two MMAs reproduce the f32 product a*s*w to ~2^-17, i.e. the
//     FP32-accumulate class of the CUDA-core GEMV (the reference's tensor-core path rounds weight*scale to BF16
//     once, a 2^-9 step).  The second MMA is free: the kernel is bound by the weight stream, not the tensor pipe.
//   * one elected thread issues 2 x 8 x tcgen05.mma.kind::f16 (128 x NB x 16, BF16 in, F32 accumulate in TMEM)
//     per k block from shared-memory descriptors, then tcgen05.commit -> mbarrier releases the stage; two
//     BF16 operand stages per CTA, one CTA per SM.
//   * epilogue: tcgen05.ld 32x32b -> registers, [split-K partials folded in split order by the last-arriving
//     CTA], bias / residual with the reference's rounding points, BF16 store (lanes = consecutive n: coalesced).
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

constexpr int kTcBK = 128;        // k block = FP8 scale block
constexpr int kTcThreads = 512;    // 16 converter warps (+ one MMA-issuing warp per CTA): the FP8 -> BF16 conversion is CUDA-core work and needs the warps to hide its latencies
constexpr int kTcMaxKb = 256;     // k blocks per CTA (K <= 32768)
constexpr int kTcStages = 2;      // BF16 operand stages (tensor-core side)
// raw FP8 weight blocks / activation slabs in flight per CTA (cp.async rings): as deep as 227 KB of shared memory
// allows next to the two BF16 operand stages -- the ring depth is what keeps HBM requests in flight while the CTA converts
__host__ __device__ constexpr int tc_raw_depth(int nb) { return nb <= 16 ? 6 : (nb <= 32 ? 5 : 3); }

struct TcParams {
    const uint16_t* A; const uint8_t* B; const uint16_t* S; uint16_t* C;
    const uint16_t* bias; const uint16_t* residual;
    int M, K, N, splits;
    float* ws; unsigned* counters;
};

// four e4m3 codes (one 32-bit word) -> four bf16 (two packed words), exact
__device__ __forceinline__ void e4m3x4_to_bf16x4(uint32_t w, uint32_t& lo, uint32_t& hi) {
    const uint32_t h0 = e4m3x2_to_f16x2_lo(w), h1 = e4m3x2_to_f16x2_hi(w);
    const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&h0));
    const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&h1));
    lo = pack_bf16x2(f0.x, f0.y);
    hi = pack_bf16x2(f1.x, f1.y);
}

template <int NB>
__global__ void __launch_bounds__(kTcThreads + 32, 1)
gemm_fp8_tc_kernel(TcParams p) {
    constexpr int NW = kTcThreads / 32;
    constexpr int kTcRaw = tc_raw_depth(NB);
    constexpr int WI = 32 / NW;                   // weight items per thread per k block (8 rows x 64 B each per warp)
    constexpr int AI = (NB / 2 + NW - 1) / NW;    // activation items per thread per k block (copy: 2 rows per warp)
    constexpr uint32_t A_STAGE = kTcRows * kTcBK * 2;      // 32 KB of BF16 weights
    constexpr uint32_t B_HALF = NB * kTcBK * 2;            // one BF16 activation tile (<= 16 KB)
    constexpr uint32_t B_STAGE = 2 * B_HALF;               // hi + lo terms of activation * block scale
    constexpr uint32_t TMEM_COLS = NB < 32 ? 32 : NB;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t mma_bar[kTcStages];
    __shared__ uint32_t tmem_base_s;
    __shared__ unsigned s_ticket;
    __shared__ float sc_s[kTcMaxKb];              // this CTA's block scales (one per k block), staged once

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, split = blockIdx.y;
    const int n0 = tile * kTcRows;
    const int nkb_total = p.K / kTcBK;
    const int per = (nkb_total + p.splits - 1) / p.splits;
    const int kb0 = split * per, kb1 = min(nkb_total, kb0 + per);
    const int nkb = max(0, kb1 - kb0);
    const int SK = nkb_total;                     // scale row length (K % 128 == 0)
    const int rr = lane & 7, piece = lane >> 3;

    // block scales are weights: staged before the dependency wait, off the per-block critical path
    for (int i = tid; i < nkb; i += blockDim.x) sc_s[i] = bf16_bits_to_f32(__ldg(p.S + (size_t)(n0 >> 7) * SK + kb0 + i));
    if (warp == 0) tmem_alloc(&tmem_base_s, TMEM_COLS);
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kTcStages; ++s) mbar_init(&mma_bar[s], 1);
        mbar_fence_init();
    }

    // ---- raw rings (cp.async / LDGSTS): global -> shared in FULL 128-byte lines (8 lanes per weight row, 16 lanes
    // per activation row), stored row-major with the 16-byte pieces of a row XOR-swizzled by (row & 7) so that
    // both the line-wise copy and the core-matrix-wise read-back (8 rows x one piece per quarter warp) are
    // bank-conflict free.  Weights never depend on the previous kernel: their first kTcRaw blocks are requested
    // before the dependency wait, the activation slabs after it.  Group order: W0..W(D-1) | A0..A(D-1) |
    // (W_D, A_D) | ...: block `b` is complete for a thread once at most D-1 of its groups are pending, and
    // visible to the CTA after the next barrier -- iteration `it` therefore waits for block it+1 before its barrier.
    unsigned char* raw_w = smem + (size_t)kTcStages * (A_STAGE + B_STAGE);
    unsigned char* raw_a = raw_w + (size_t)kTcRaw * kTcRows * kTcBK;
    auto issue_weights = [&](int kb, int rs) {
#pragma unroll
        for (int i = 0; i < WI; ++i) {
            const int r = (i * NW + warp) * 4 + (lane >> 3), pc = lane & 7;     // 4 rows x 128 B per warp request
            const uint8_t* src = p.B + (size_t)(n0 + r) * p.K + (size_t)kb * kTcBK + pc * 16;
            const uint32_t dst = smem_u32(raw_w + (size_t)rs * (kTcRows * kTcBK) + r * 128 + ((pc ^ (r & 7)) * 16));
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
        }
    };
    auto issue_acts = [&](int kb, int rs) {
#pragma unroll
        for (int i = 0; i < AI; ++i) {
            const int m = (i * NW + warp) * 2 + (lane >> 4), pc = lane & 15;    // 2 rows x 256 B per warp request
            if (m < NB) {
                const bool ok = m < p.M;                       // rows past the batch are zero-filled (src-size 0)
                const uint16_t* src = p.A + (ok ? (size_t)m * p.K + (size_t)kb * kTcBK + pc * 8 : 0);
                const uint32_t dst = smem_u32(raw_a + (size_t)rs * (NB * kTcBK * 2) + m * 256 + ((pc ^ (m & 7)) * 16));
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
            }
        }
    };
#pragma unroll
    const bool conv = warp < NW;                  // 16 converter warps; warp NW only issues the MMAs
    for (int d = 0; d < kTcRaw; ++d) {
        if (conv && d < nkb) issue_weights(kb0 + d, d);
        asm volatile("cp.async.commit_group;" ::: "memory");   // one group per block, empty ones included
    }
    griddep_wait();
    griddep_launch();
#pragma unroll
    for (int d = 0; d < kTcRaw; ++d) {
        if (conv && d < nkb) issue_acts(kb0 + d, d);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.wait_group %0;" :: "n"(kTcRaw - 1) : "memory");       // block 0 (weights + slab) landed

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_s;
    constexpr uint32_t idesc = umma_idesc_bf16(NB);

    for (int it = 0; it < nkb; ++it) {
        const int st = it & 1, kb = kb0 + it, rs = it % kTcRaw;
        unsigned char* a_s = smem + (size_t)st * (A_STAGE + B_STAGE);
        unsigned char* b_s = a_s + A_STAGE;
        const unsigned char* rw = raw_w + (size_t)rs * (kTcRows * kTcBK);
        const unsigned char* ra = raw_a + (size_t)rs * (NB * kTcBK * 2);
        const float sc = sc_s[it];
        if (conv) {
            uint4 wq[WI], aq[AI];
#pragma unroll
            for (int i = 0; i < WI; ++i) {
                const int q = i * NW + warp;                      // 32 items: row group q>>1, 64-byte half q&1
                const int r = 8 * (q >> 1) + rr, pc = (q & 1) * 4 + piece;
                wq[i] = *reinterpret_cast<const uint4*>(rw + r * 128 + ((pc ^ rr) * 16));
            }
#pragma unroll
            for (int i = 0; i < AI; ++i) {
                const int q = i * NW + warp;                      // NB/2 items: row group q>>2, 64-byte quarter q&3
                if (q < NB / 2) {
                    const int m = 8 * (q >> 2) + rr, pc = (q & 3) * 4 + piece;
                    aq[i] = *reinterpret_cast<const uint4*>(ra + m * 256 + ((pc ^ rr) * 16));
                }
            }
            // the MMAs that read this BF16 stage two iterations ago must have completed
            if (it >= kTcStages) mbar_wait(&mma_bar[st], (uint32_t)(((it >> 1) - 1) & 1));
            // weights: 16 codes -> 16 bf16 = k chunks c0, c0 + 1 of row r
#pragma unroll
            for (int i = 0; i < WI; ++i) {
                const int q = i * NW + warp;
                const int c0 = (q & 1) * 8 + piece * 2;
                unsigned char* dst = a_s + (size_t)c0 * (kTcRows * 16) + (size_t)(q >> 1) * 128 + rr * 16;
                uint4 o0, o1;
                e4m3x4_to_bf16x4(wq[i].x, o0.x, o0.y); e4m3x4_to_bf16x4(wq[i].y, o0.z, o0.w);
                e4m3x4_to_bf16x4(wq[i].z, o1.x, o1.y); e4m3x4_to_bf16x4(wq[i].w, o1.z, o1.w);
                *reinterpret_cast<uint4*>(dst) = o0;
                *reinterpret_cast<uint4*>(dst + kTcRows * 16) = o1;
            }
            // activations: 8 bf16 x block scale -> bf16, k chunk c of batch row m
#pragma unroll
            for (int i = 0; i < AI; ++i) {
                const int q = i * NW + warp;
                if (q < NB / 2) {
                    const int c = (q & 3) * 4 + piece;
                    unsigned char* dst = b_s + (size_t)c * (NB * 16) + (size_t)(q >> 2) * 128 + rr * 16;
                    const uint32_t aw[4] = {aq[i].x, aq[i].y, aq[i].z, aq[i].w};
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float v0 = bf16lo(aw[j]) * sc, v1 = bf16hi(aw[j]) * sc;   // exact-ish f32 (8 x 8 bit product)
                        hi[j] = pack_bf16x2(v0, v1);
                        lo[j] = pack_bf16x2(v0 - bf16lo(hi[j]), v1 - bf16hi(hi[j]));    // residual of the BF16 rounding
                    }
                    *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    *reinterpret_cast<uint4*>(dst + B_HALF) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                }
            }
            fence_proxy_async();          // generic-proxy smem writes -> visible to the tensor core (async proxy)
            asm volatile("cp.async.wait_group %0;" :: "n"(kTcRaw - 2) : "memory");   // this thread's part of block it+1 landed
        }
        __syncthreads();                  // BF16 stage complete; raw slot `rs` fully read; block it+1 visible to all
        if (conv) {
            // refill the raw slot that every thread has finished reading, keep the group count uniform
            if (it + kTcRaw < nkb) { issue_weights(kb + kTcRaw, rs); issue_acts(kb + kTcRaw, rs); }
            asm volatile("cp.async.commit_group;" ::: "memory");
        } else if (lane == 0) {           // the MMA warp: off the converters' critical path
            tc_fence_after();
            const uint32_t a_addr = smem_u32(a_s), b_addr = smem_u32(b_s);
#pragma unroll
            for (int ks = 0; ks < kTcBK / 16; ++ks) {
                const uint64_t da = umma_desc(a_addr + ks * 2 * (kTcRows * 16), kTcRows * 16, 128);
                const uint64_t db = umma_desc(b_addr + ks * 2 * (NB * 16), NB * 16, 128);
                const uint64_t dl = umma_desc(b_addr + B_HALF + ks * 2 * (NB * 16), NB * 16, 128);
                umma_bf16(tmem_d, da, db, idesc, (it > 0 || ks > 0) ? 1u : 0u);
                umma_bf16(tmem_d, da, dl, idesc, 1u);
            }
            umma_commit(&mma_bar[st]);    // arrives when every MMA issued so far has completed
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ---- epilogue
    if (nkb > 0) {
        const int last = nkb - 1;
        mbar_wait(&mma_bar[last & 1], (uint32_t)((last >> 1) & 1));
    }
    tc_fence_after();
    const int row = 32 * (warp & 3) + lane;                 // TMEM lane = weight row of the tile
    const int n = n0 + row;
    const size_t tile_ws = ((size_t)tile * p.splits + split) * (size_t)NB * kTcRows;
    if (warp < 4) {
#pragma unroll
        for (int c = 0; c < NB; c += 16) {
            float v[16];
            if (nkb > 0) tmem_ld16(tmem_d + ((uint32_t)(32 * warp) << 16) + c, v);
            else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int m = c + j;
                if (m >= p.M) continue;
                if (p.splits == 1) {
                    float x = v[j];
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
    // last CTA of the tile: fold the split-K partials in split order (deterministic), then the epilogue.
    // thread = (weight row, half of the batch rows); the loads of eight splits are issued together.
    if (conv) {
        const int r = tid & (kTcRows - 1), half = tid >> 7;      // four batch-row phases with 512 threads
        const float* base = p.ws + (size_t)tile * p.splits * NB * kTcRows;
        const int nn = n0 + r;
        for (int m = half; m < p.M; m += kTcThreads / kTcRows) {
            float x = 0.f;
            for (int s0 = 0; s0 < p.splits; s0 += 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    v[u] = (s0 + u < p.splits) ? __ldcg(base + ((size_t)(s0 + u) * NB + m) * kTcRows + r) : 0.f;
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

static float* g_tc_ws = nullptr;
static unsigned* g_tc_cnt = nullptr;
static const size_t kTcWsBytes = 96ull << 20;
static const int kTcMaxTiles = 8192;

template <int NB>
static int launch_tc(TcParams p, cudaStream_t st) {
    const int tiles = p.N / kTcRows, nkb = p.K / kTcBK;
    // one CTA per SM: pick the split-K count that fills whole waves of the 148 SMs best (ties -> fewer splits,
    // i.e. fewer partials through L2); every split keeps at least two k blocks
    const int sms = sm_count();
    int splits = 1;
    double best = -1.0;
    for (int s = 1; s <= 12 && s * 2 <= (nkb < 2 ? 2 : nkb); ++s) {
        const long ctas = (long)tiles * s;
        const long waves = (ctas + sms - 1) / sms;
        const int per = (nkb + s - 1) / s;
        // time ~ waves * (k blocks per CTA + fixed cost of ~2 blocks): lower is better
        const double t = (double)waves * (per + 2.0);
        const double score = 1.0 / t - 1e-6 * s;
        if (score > best) { best = score; splits = s; }
    }
    while (splits > 1 && (size_t)tiles * splits * NB * kTcRows * sizeof(float) > kTcWsBytes) --splits;
    if (!g_tc_ws) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(st, &cs);
        PGK_REQUIRE(cs == cudaStreamCaptureStatusNone,
                    "gemm_fp8_tc failed: first call must happen outside graph capture (workspace allocation)");
        PGK_CUDA(cudaMalloc(&g_tc_ws, kTcWsBytes));
        PGK_CUDA(cudaMalloc(&g_tc_cnt, kTcMaxTiles * sizeof(unsigned)));
        PGK_CUDA(cudaMemset(g_tc_cnt, 0, kTcMaxTiles * sizeof(unsigned)));
    }
    p.splits = splits;
    p.ws = g_tc_ws;
    p.counters = g_tc_cnt;
    constexpr size_t kSmem = (size_t)kTcStages * (kTcRows * kTcBK * 2 + 2 * NB * kTcBK * 2) +
                             (size_t)tc_raw_depth(NB) * (kTcRows * kTcBK + NB * kTcBK * 2);
    static bool attr_set = false;
    if (!attr_set) {
        PGK_CUDA(cudaFuncSetAttribute(gemm_fp8_tc_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
        attr_set = true;
    }
    PGK_CUDA(launch(gemm_fp8_tc_kernel<NB>, dim3(tiles, splits), dim3(kTcThreads + 32), kSmem, st, p));
    return 0;
}

bool gemm_fp8_tc_shape_ok(int M, int K, int N, const void* A, const void* B);

// Smallest batch sent to the tensor cores.  The kernel zero-fills rows M..NB-1, so any M >= 1 works; the crossover
// against the CUDA-core slab GEMV (whose time grows linearly with M: dequant + M FMAs per weight) measured on
// B200 is between M = 4 (15.9 us GEMV vs ~16 us here, qkv shape) and M = 5.  PGK_TC_MIN_M overrides (A/B timing).
int gemm_fp8_tc_min_m() {
    static const int v = [] { const char* e = getenv("PGK_TC_MIN_M"); int m = e ? atoi(e) : 5; return m < 1 ? 1 : m; }();
    return v;
}

// the dispatcher's test (batch >= crossover); the kernel itself takes any 1 <= M <= 64
bool gemm_fp8_tc_eligible(int M, int K, int N, const void* A, const void* B) {
    return M >= gemm_fp8_tc_min_m() && gemm_fp8_tc_shape_ok(M, K, N, A, B);
}

bool gemm_fp8_tc_shape_ok(int M, int K, int N, const void* A, const void* B) {
    return M >= 1 && M <= 64 && K / kTcBK <= kTcMaxKb && K % kTcBK == 0 && N % kTcRows == 0 && N / kTcRows <= kTcMaxTiles &&
           ((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0);
}

int gemm_fp8_tc(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N, const void* bias,
                const void* residual, cudaStream_t st) {
    PGK_REQUIRE(gemm_fp8_tc_shape_ok(M, K, N, A, B_nk),
                "gemm_fp8_tc failed: needs 1 <= M <= 64, K %% 128 == 0, N %% 128 == 0 and 16-byte aligned pointers "
                "(M=%d K=%d N=%d)", M, K, N);
    TcParams p = {};
    p.A = (const uint16_t*)A; p.B = (const uint8_t*)B_nk; p.S = (const uint16_t*)S; p.C = (uint16_t*)C;
    p.bias = (const uint16_t*)bias; p.residual = (const uint16_t*)residual;
    p.M = M; p.K = K; p.N = N;
    if (M <= 16) return launch_tc<16>(p, st);
    if (M <= 32) return launch_tc<32>(p, st);
    return launch_tc<64>(p, st);
}

}  // namespace pgk

extern "C" int pgk_gemm_fp8_bf16_tc(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N,
                                    const void* bias, const void* residual, void* stream) {
    bool clean = true;
    if (int rc = pgk::fp8_codes_clean(B_nk, (size_t)N * K, (cudaStream_t)stream, &clean)) return rc;
    PGK_REQUIRE(clean, "gemm_fp8_bf16_tc failed: the weights hold E4M3 codes 0x7F/0xFF (+-480 in the reference's tables, "
                "NaN for the tensor-core conversion); use pgk_gemv_fp8_bf16, which serves them exactly");
    return pgk::gemm_fp8_tc(A, B_nk, S, C, M, K, N, bias, residual, (cudaStream_t)stream);
}
