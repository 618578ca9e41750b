// step.cu -- the one-launch decode step: a persistent kernel that runs a whole token (every layer's fused
// RMSNorm + W8A16/NVF4 GEMV, QK-norm + RoPE + KV write + split-KV attention, SwiGLU, residual GEMVs, final norm +
// lm_head + greedy argmax) as a "step program" of ops with device-side grid barriers between them.
//
// replaces the launch sequence of DecodeM1.step (src/pygpukit/llm/decode/m1.py:40-121) and, on this repo's own
// path, the 5-launches-per-layer graph of llm/models/causal.py::decode_step_device.
//
// Design (HBM-bound; DESIGN.md "step program"):
//   * grid = one CTA per SM, 16 consumer warps + 1 producer warp.  The producer walks the SAME op list as the
//     consumers and streams every byte the CTA will need (weight rows, K/V tiles) into a ring of 8 KB shared-memory
//     slots with cp.async.bulk (TMA, SASS UBLKCP) completing on per-slot mbarriers.  Weights and old KV rows are
//     constants of the step, so the producer never waits for a grid barrier: while the consumers synchronise and
//     re-stage the activation vector (~2 us per op boundary) the ring keeps filling (up to ~190 KB per SM = 28 MB
//     in flight on the GPU), and the HBM stream does not stop at op boundaries.
//   * work split: an op's rows are cut into items (the rows that fit one slot, or one row cut into <= 4 slot-sized
//     segments), CTAs own contiguous item ranges (= contiguous byte ranges of the weight matrix), items are dealt
//     round-robin to the consumer warps.  One warp owns a row, lane l owns the 32-byte chunks l, l+32, ... in
//     increasing order -- exactly the per-row order of gemv_fp8_sb_kernel / gemv_bf16_kernel, so results are
//     bit-identical to the per-op entry points.
//   * attention: units of 16 KV positions (4 KB of K + 4 KB of V = one slot) over the flattened (kv head, tile)
//     space, contiguous unit ranges per CTA, one warp per unit with the full online-softmax state of the G <= 8
//     query heads of the group in registers (lane = 4 head dims); per-(kv head, CTA) partial records in L2, then a
//     grid barrier and a distributed fold in CTA order (deterministic).
//   * grid barrier: one monotone counter per step parity (red.release.gpu + ld.acquire.gpu spin by one thread,
//     named barrier for the consumer threads).  Every wait is bounded: on a timeout the kernel records an error
//     code and runs to completion without waiting any more (results are then garbage; pgk_step_error reports it).
#include "common.cuh"
#include "gemv_common.cuh"
#include "../../include/pgk_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace pgk {

// p2p.cu: LL regions of the peer-memory buffers (tensor parallelism)
bool p2p_ll_info(int* rank, int* world, unsigned char** ll_peers, unsigned** epoch, size_t* words_per_source);

constexpr int kCW = 14;                 // consumer warps
constexpr int kNP = 2;                  // producer warps (one issuing thread each; slot q belongs to producer q % kNP)
constexpr int kCT = kCW * 32;           // consumer threads
constexpr int kStepThreads = kCT + 32 * kNP;   // 512 threads: 128 registers each
constexpr int kSlotB = 16384;           // ring slot bytes
constexpr int kMaxSlots = 13;
// "full" barriers per slot, used round-robin by ring round.  A parity wait is only meaningful when the waiter is at
// most one phase ahead of the barrier; a consumer warp jumps kCW items ahead after each item, i.e. up to
// (kCW - 1) * spi + 1 slots beyond the fills it has seen complete, which can exceed one ring round when an item spans
// several slots.  With kPhases barriers per slot a warp would have to run kPhases rounds ahead before a wait could be
// misread (pgk_step_create checks kPhases * slots >= (kCW - 1) * spi + 1).
constexpr int kPhases = 8;
constexpr int kAttnT = 32;              // KV positions per attention unit (K tile 8 KB + V tile 8 KB = one slot)
constexpr int kRec = 130;               // partial record: o[128], m, l
constexpr int kPartStride = 132;        // shared-memory copy of a record (float4 aligned)
constexpr int kTracePts = 8;            // trace stamps per op (CTA 0)
constexpr long long kSpinLimit = 3000000000LL;   // cycles (~1.5 s): far beyond any legitimate wait
constexpr float kNeg = -1e30f;

enum : int { OP_FP8 = 1, OP_BF16 = 2, OP_NVF4R = 3, OP_ATTN = 4 };
enum : int { F_SWIGLU = 1, F_ACT_EMBED = 2, F_RES_EMBED = 4, F_ARGMAX = 8, F_XCHG = 16 };
constexpr int kXMax = 128;              // rows of a row-parallel op one CTA may own under tensor parallelism
constexpr int kMaxWorld = 8;
constexpr int kScalePad = 256;           // k blocks per staged scale row (K <= 32768)

struct StepHeader {
    const pgk_step_op_t* ops;
    int n_ops, ns, act_bytes, vocab, advance;
    int bar_mode;                  // grid-barrier variant (A/B switch, PGK_STEP_BAR)
    int inflight;                  // fills a producer may have outstanding (issued, not landed): bounds the memory queue
    int32_t* token_buf; int32_t* pos_buf; int32_t* ctx_buf; int32_t* next_token;
    unsigned* bar;                 // [0],[1] grid-barrier counters by step parity, [2] step counter
    float* attn_ws;                // [ctas][RT][kRec]
    unsigned long long* am_key;    // argmax scratch (zero between steps)
    int* error;
    unsigned long long* trace;     // [(n_ops + 1) * kTracePts] globaltimer (ns) stamps of CTA 0
    // tensor parallelism (tp_world == 1: unused).  ll[r] = the LL region of rank r as mapped here; per parity and
    // source rank `ll_words` words of 8 bytes {payload, exchange number}
    int tp_rank, tp_world, vocab_offset, n_xchg;
    unsigned char* ll[kMaxWorld];
    unsigned* ll_epoch;            // device word: exchanges completed so far (identical on every rank)
    unsigned ll_words;
};

// ---------------------------------------------------------------- work split, shared by host validation, the
// producers and the consumers (all must agree exactly).  32-bit arithmetic only: 64-bit division is a ~100-instruction
// routine on the GPU.
struct GemvPlan {
    int RB, nmat, R, nseg, seg_bytes, spi, items_total, item0, n_items;
};
__host__ __device__ inline int op_row_bytes(int kind, int K) {
    return kind == OP_FP8 ? K : (kind == OP_BF16 ? 2 * K : (K / 2 + K / 32));
}
__host__ __device__ inline unsigned part_lo(unsigned total, unsigned cta, unsigned ncta) { return total * cta / ncta; }
__host__ __device__ inline GemvPlan gemv_plan(int kind, int flags, int K, int N, int cta, int ncta) {
    GemvPlan p;
    p.RB = op_row_bytes(kind, K);
    p.nmat = (flags & F_SWIGLU) ? 2 : 1;
    if (p.nmat * p.RB <= kSlotB) {
        p.R = kSlotB / (p.nmat * p.RB); p.nseg = 1; p.seg_bytes = p.RB; p.spi = 1;
    } else {
        p.R = 1; p.nseg = (p.RB + kSlotB - 1) / kSlotB; p.seg_bytes = p.RB / p.nseg; p.spi = p.nmat * p.nseg;
    }
    p.items_total = (N + p.R - 1) / p.R;
    p.item0 = (int)part_lo((unsigned)p.items_total, (unsigned)cta, (unsigned)ncta);
    p.n_items = (int)part_lo((unsigned)p.items_total, (unsigned)cta + 1u, (unsigned)ncta) - p.item0;
    return p;
}
// attention: the CTAs are dealt to the KV heads in contiguous groups (CTA c serves head c * n_kv / ncta), the head's
// tiles of kAttnT positions are dealt evenly over the CTAs of its group
constexpr int kAttnMaxGroup = 48;       // CTAs that share one KV head (= partial records the fold reads per head)
struct AttnPlan {
    int kvh, cfirst, gs, ntile, u0, u1;
};
__host__ __device__ inline int attn_ctx(int ctx, int max_seq) { return ctx < 1 ? 1 : (ctx > max_seq ? max_seq : ctx); }
__host__ __device__ inline int attn_cfirst(int k, int n_kv, int ncta) { return (k * ncta + n_kv - 1) / n_kv; }
__host__ __device__ inline AttnPlan attn_plan(int n_kv, int ctx, int cta, int ncta) {
    AttnPlan a;
    a.kvh = cta * n_kv / ncta;
    a.cfirst = attn_cfirst(a.kvh, n_kv, ncta);
    a.gs = attn_cfirst(a.kvh + 1, n_kv, ncta) - a.cfirst;
    if (a.gs > kAttnMaxGroup) a.gs = kAttnMaxGroup;      // the other CTAs of the group sit this op out
    a.ntile = (ctx + kAttnT - 1) / kAttnT;
    const int lc = cta - a.cfirst;
    a.u0 = lc < a.gs ? lc * a.ntile / a.gs : 0;
    a.u1 = lc < a.gs ? (lc + 1) * a.ntile / a.gs : 0;
    return a;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------- bounded waits
struct Abort {
    volatile int* s_abort;   // shared: this CTA gave up waiting
    int* gerr;               // global error word
};
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __noinline__ void mbar_wait_slow(uint64_t* bar, uint32_t parity, Abort ab, int code) {
    const long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        if (*ab.s_abort) return;
        if (clock64() - t0 > kSpinLimit) { *ab.s_abort = 1; atomicCAS(ab.gerr, 0, code); return; }
    }
}
__device__ __forceinline__ void mbar_wait_b(uint64_t* bar, uint32_t parity, Abort ab, int code) {
    if (mbar_try(bar, parity)) return;
    mbar_wait_slow(bar, parity, ab, code);
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_inc(unsigned* p) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(p) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" :: "n"(kCT) : "memory"); }

// all consumer threads of every CTA: arrive + wait on the step's monotone counter.  The CTA barrier orders every
// consumer thread's global writes before thread 0's release (cumulativity); the acquire + second CTA barrier order the
// other CTAs' writes before every thread's later loads (which bypass L1: ld.global.cg / ld.global.nc of constants only).
__device__ __forceinline__ unsigned ld_relaxed_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// mode 0: red.release + ld.acquire polls.  mode 1: fence, relaxed atomic, relaxed polls, fence.  mode 2: as 1 with a
// 64 ns back-off between polls.  mode 3: as 1, the atomic returns (its round trip is the first poll).
__device__ __forceinline__ void grid_sync(unsigned* ctr, unsigned target, int tid, Abort ab, unsigned long long* tr, int mode) {
    consumer_sync();
    if (tid == 0) {
        if (tr) tr[0] = globaltimer_ns();
        unsigned seen = 0;
        if (mode == 0) {
            red_release_gpu_inc(ctr);
            seen = ld_acquire_gpu(ctr);
        } else {
            __threadfence();
            seen = atomicAdd(ctr, 1u) + 1u;
        }
        if ((int)(seen - target) < 0) {
            const long long t0 = clock64();
            while (true) {
                seen = mode == 0 ? ld_acquire_gpu(ctr) : ld_relaxed_gpu(ctr);
                if ((int)(seen - target) >= 0) break;
                if (*ab.s_abort) break;
                if (clock64() - t0 > kSpinLimit) { *ab.s_abort = 1; atomicCAS(ab.gerr, 0, 100); break; }
                if (mode == 2) __nanosleep(64);
            }
        }
        if (mode != 0) __threadfence();
    }
    consumer_sync();
}

__device__ __forceinline__ uint64_t make_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_copy_hint(uint32_t smem_dst, const void* gmem_src, uint32_t bytes, uint32_t bar,
                                               uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(smem_dst), "l"(gmem_src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_copy_plain(uint32_t smem_dst, const void* gmem_src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_dst), "l"(gmem_src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_u32(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}

// this lane's 32-byte chunk (lane stride 32 bytes) without bank conflicts: the lanes of a quarter warp read the
// two 16-byte halves in opposite order (lanes 0-3 low half first, lanes 4-7 high half first, ...), so the eight
// 16-byte requests of a phase fall into eight different bank groups; the halves are swapped back in registers
__device__ __forceinline__ u32x8 lds_chunk32(const unsigned char* p, int lane) {
    const int sw = (lane >> 2) & 1;
    const uint4 a = *reinterpret_cast<const uint4*>(p + 16 * sw);
    const uint4 b = *reinterpret_cast<const uint4*>(p + 16 * (sw ^ 1));
    u32x8 r;
    r.v[0] = sw ? b.x : a.x; r.v[1] = sw ? b.y : a.y; r.v[2] = sw ? b.z : a.z; r.v[3] = sw ? b.w : a.w;
    r.v[4] = sw ? a.x : b.x; r.v[5] = sw ? a.y : b.y; r.v[6] = sw ? a.z : b.z; r.v[7] = sw ? a.w : b.w;
    return r;
}

// where the bytes of an item live in the ring
struct ItemView {
    const unsigned char* ring;
    int ns, qmod;                 // ring slots, first slot of the item (already reduced mod ns)
    int R, RB, nseg, seg_bytes, spi;
    __device__ __forceinline__ const unsigned char* ptr(int t, int r, int off) const {
        if (spi == 1) return ring + (size_t)qmod * kSlotB + (size_t)(t * R + r) * RB + off;
        const int seg = (off >= seg_bytes) + (off >= 2 * seg_bytes) + (off >= 3 * seg_bytes);
        int idx = qmod + t * nseg + seg;
        if (idx >= ns) idx -= ns;
        return ring + (size_t)idx * kSlotB + (off - seg * seg_bytes);
    }
};

// ---------------------------------------------------------------- activation staging by the consumer threads
// Same arithmetic as stage_activations_f16 / stage_activations_bf16 (gemv_common.cuh, gemv_bf16.cu): the row is read
// once, the RMSNorm statistics follow the reference's summation tree (row_inv_rms_ref_order), elements are
// normalised as bf16(x * inv * gamma), then either kept as bf16 (TO_F16 = false) or range-shifted to f16.
// Vector j (8 elements) lands in plane j % PL at chunk j / PL.   red: 80 words of shared scratch.
// FP8 block scales of the rows a CTA owns (at most two 128-row blocks, one or two matrices): loaded together with the
// activation vector (one L2 round trip for both) and parked in shared memory for the row loop
struct ScaleJob {
    const uint16_t* S0; const uint16_t* S1;
    int count, SK, rb0, nrb;       // count = matrices * 2 * SK values (0 = no job)
    float* dst;                    // [matrix][row block 0/1][kScalePad]
};
template <int PL, bool TO_F16, int VY>
__device__ __forceinline__ void stage_act_v(const uint16_t* __restrict__ A, int K, const uint16_t* __restrict__ gamma,
                                            float eps, unsigned char* act, uint32_t* red, float* inv_s, int tid,
                                            const ScaleJob& sj, unsigned long long* tr) {
    const int nvec = K >> 3, lane = tid & 31, warp = tid >> 5;
    const int nchunk = nvec / PL;
    const uint4* src = reinterpret_cast<const uint4*>(A);
    const uint4* gv = reinterpret_cast<const uint4*>(gamma);
    uint4* dst = reinterpret_cast<uint4*>(act);
    float* red_f = reinterpret_cast<float*>(red) + 64;
    constexpr bool EARLY_G = VY <= 4;      // gamma in flight together with the activations (one L2 round trip, not two)
    uint4 v[VY], g[EARLY_G ? VY : 1];
#pragma unroll
    for (int e = 0; e < VY; ++e) {
        const int j = tid + e * kCT;
        if (j < nvec) {
            v[e] = ld_act16(src + j);
            if (EARLY_G && gamma) g[e] = __ldg(gv + j);
        }
    }
    float scv[2] = {0.f, 0.f};
    int sci[2] = {-1, -1};
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int i = tid + u * kCT;
        if (i < sj.count) {
            const int t = i / (2 * sj.SK), rem = i - t * 2 * sj.SK, b = rem / sj.SK, kb = rem - b * sj.SK;
            sci[u] = (t * 2 + b) * kScalePad + kb;
            if (sj.rb0 + b < sj.nrb) scv[u] = bf16_bits_to_f32(__ldg((t ? sj.S1 : sj.S0) + (size_t)(sj.rb0 + b) * sj.SK + kb));
        }
    }
    if (gamma) {
#pragma unroll
        for (int e = 0; e < VY; ++e) {
            const int j = tid + e * kCT;
            if (j < nvec) dst[j] = v[e];
        }
        consumer_sync();
        if (tr && tid == 0) tr[5] = globaltimer_ns();
        const uint16_t* raw = reinterpret_cast<const uint16_t*>(dst);
        const int bs = min(256, (K + 31) & ~31), nvw = bs >> 5;
        for (int w = warp; w < nvw; w += kCW) {
            float part = 0.f;
            for (int i = w * 32 + lane; i < K; i += bs) { float x = bf16_bits_to_f32(raw[i]); part = fmaf(x, x, part); }
            part = warp_sum(part);
            if (lane == 0) red_f[w] = part;
        }
        consumer_sync();
        if (tr && tid == 0) tr[6] = globaltimer_ns();
        const float ss = warp_sum(lane < nvw ? red_f[lane] : 0.f);
        const float rinv = rsqrtf(ss / (float)K + eps);
#pragma unroll
        for (int e = 0; e < VY; ++e)
            if (tid + e * kCT < nvec) v[e] = norm8_bf16(v[e], EARLY_G ? g[e] : __ldg(gv + tid + e * kCT), rinv);
    }
    float mult = 1.f;
    if (TO_F16) {
        uint32_t mx = 0;
#pragma unroll
        for (int e = 0; e < VY; ++e) {
            if (tid + e * kCT < nvec) {
                uint32_t t = __vmaxu2(__vmaxu2(v[e].x & 0x7fff7fffu, v[e].y & 0x7fff7fffu),
                                      __vmaxu2(v[e].z & 0x7fff7fffu, v[e].w & 0x7fff7fffu));
                mx = max(mx, max(t & 0xffffu, t >> 16));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) red[warp] = mx;
        consumer_sync();                 // also: every thread is done reading the raw row
        if (tr && tid == 0) tr[7] = globaltimer_ns();
        uint32_t bm = 0;
        for (int w = 0; w < kCW; ++w) bm = max(bm, red[w]);
        int ex = (int)(bm >> 7);
        ex = min(max(ex, 15), 254);
        if ((bm >> 7) == 255) ex = 141;
        mult = __uint_as_float((uint32_t)(268 - ex) << 23);
        if (tid == 0) inv_s[0] = __uint_as_float((uint32_t)(ex - 14) << 23);
    } else {
        consumer_sync();                 // every thread is done reading the raw row
    }
#pragma unroll
    for (int e = 0; e < VY; ++e) {
        const int j = tid + e * kCT;
        if (j < nvec) dst[(j % PL) * nchunk + (j / PL)] = TO_F16 ? bf16x8_to_f16x8_scaled(v[e], mult) : v[e];
    }
#pragma unroll
    for (int u = 0; u < 2; ++u)
        if (sci[u] >= 0) sj.dst[sci[u]] = scv[u];
    consumer_sync();
}
// four vectors per thread for K <= 14336 (no spills), eight for the long rows (K up to 28672)
template <int PL, bool TO_F16>
__device__ __forceinline__ void stage_act(const uint16_t* __restrict__ A, int K, const uint16_t* __restrict__ gamma,
                                          float eps, unsigned char* act, uint32_t* red, float* inv_s, int tid,
                                          const ScaleJob& sj, unsigned long long* tr) {
    if ((K >> 3) <= 4 * kCT) stage_act_v<PL, TO_F16, 4>(A, K, gamma, eps, act, red, inv_s, tid, sj, tr);
    else                     stage_act_v<PL, TO_F16, 8>(A, K, gamma, eps, act, red, inv_s, tid, sj, tr);
}

// ---------------------------------------------------------------- row dot products
__device__ __forceinline__ float dot32_e4m3_regs(const u32x8& w, const uint4 (&a)[4]) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int p = 0; p < 4; ++p) dot8_e4m3(s0, s1, w.v[2 * p], w.v[2 * p + 1], a[p]);
    return s0 + s1;
}
// 8 e2m1 weights (one 32-bit word: byte b = k 2b (low nibble), 2b+1) x 8 staged f16 activations
__device__ __forceinline__ void dot8_e2m1(float& s0, float& s1, uint32_t w, const uint4& a) {
    uint32_t h;
    h = e2m1x2_to_f16x2<0>(w); fhfma_f16_lo(s0, h, a.x); fhfma_f16_hi(s1, h, a.x);
    h = e2m1x2_to_f16x2<1>(w); fhfma_f16_lo(s0, h, a.y); fhfma_f16_hi(s1, h, a.y);
    h = e2m1x2_to_f16x2<2>(w); fhfma_f16_lo(s0, h, a.z); fhfma_f16_hi(s1, h, a.z);
    h = e2m1x2_to_f16x2<3>(w); fhfma_f16_lo(s0, h, a.w); fhfma_f16_hi(s1, h, a.w);
}

struct RowOut {
    uint16_t* C; const uint16_t* residual; float inv, alpha; int flags;
    uint16_t* xout; int xlo;     // tensor-parallel partial sums: row n goes to xout[n - xlo] (shared memory) instead of C
};
// lane 0: one finished output row (the epilogue of gemv_fp8_sb_kernel / gemv_bf16_kernel)
// `res` = the residual of row n, requested by the caller BEFORE the dot product (an L2 round trip that would otherwise
// sit at the end of every row)
__device__ __forceinline__ uint16_t finish_row(const RowOut& ro, int n, float v, uint16_t res) {
    if (ro.xout) { const uint16_t pb = f32_to_bf16_bits(v); ro.xout[n - ro.xlo] = pb; return pb; }
    const uint16_t ob = ro.residual ? hadd_bf16_bits(v, res) : f32_to_bf16_bits(v);
    if (ro.C) ro.C[n] = ob;
    return ob;
}
__device__ __forceinline__ void finish_swiglu(const RowOut& ro, int n, float vg, float vu) {
    const float g = round_bf16(vg), u = round_bf16(vu);
    const float s = round_bf16(silu_ref(g));
    ro.C[n] = f32_to_bf16_bits(s * u);
}

// FP8: RR rows of NMAT matrices share one read of the activation chunk.  Per row: acc = fma(dot32(chunk), scale, acc)
// over the lane's chunks in increasing order, then the butterfly sum (= gemv_fp8_sb_kernel).
template <int NMAT, int RR>
__device__ __forceinline__ void fp8_rows(const float* sc_s, int sk_pad, int rb0, const ItemView& iv, const uint4* a_lo, int nchunk,
                                         int n, int r0, int lane, const RowOut& ro) {
    uint16_t res[RR];
#pragma unroll
    for (int rr = 0; rr < RR; ++rr) res[rr] = (lane == 0 && ro.residual && !ro.xout) ? ld_act2(ro.residual + n + rr) : (uint16_t)0;
    float acc[NMAT][RR];
#pragma unroll
    for (int t = 0; t < NMAT; ++t)
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) acc[t][rr] = 0.f;
#pragma unroll 2
    for (int c = lane; c < nchunk; c += 32) {
        uint4 a[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) a[p] = a_lo[p * nchunk + c];
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) {
#pragma unroll
            for (int t = 0; t < NMAT; ++t) {
                const u32x8 w = lds_chunk32(iv.ptr(t, r0 + rr, c * 32), lane);
                // block scales of the CTA's (at most two) 128-row blocks, staged in shared memory: [matrix][row block][k block]
                const float sc = sc_s[(t * 2 + (((n + rr) >> 7) - rb0)) * sk_pad + (c >> 2)];
                acc[t][rr] = fmaf(dot32_e4m3_regs(w, a), sc, acc[t][rr]);
            }
        }
    }
#pragma unroll
    for (int t = 0; t < NMAT; ++t)
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) acc[t][rr] = warp_sum(acc[t][rr]);
    if (lane == 0) {
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) {
            if (NMAT == 2) finish_swiglu(ro, n + rr, acc[0][rr] * ro.inv, acc[NMAT - 1][rr] * ro.inv);
            else finish_row(ro, n + rr, acc[0][rr] * ro.inv, res[rr]);
        }
    }
}

// NVF4 over row-packed weights: chunk = 32 data bytes = 64 weights = two scale blocks; per block the f16 products are
// summed in f32 and scaled by the UE4M3 block scale (exact decode), acc = fma(blocksum, scale, acc)
template <int NMAT, int RR>
__device__ __forceinline__ void nvf4_rows(const pgk_step_op_t& op, const ItemView& iv, const uint4* a_lo, int nchunk,
                                          int n, int r0, int lane, const RowOut& ro) {
    const int soff = op.K >> 1;       // scale bytes follow the data bytes of the row
    uint16_t res[RR];
#pragma unroll
    for (int rr = 0; rr < RR; ++rr) res[rr] = (lane == 0 && ro.residual && !ro.xout) ? ld_act2(ro.residual + n + rr) : (uint16_t)0;
    float acc[NMAT][RR];
#pragma unroll
    for (int t = 0; t < NMAT; ++t)
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) acc[t][rr] = 0.f;
#pragma unroll 1
    for (int c = lane; c < nchunk; c += 32) {
        uint4 a[8];
#pragma unroll
        for (int p = 0; p < 8; ++p) a[p] = a_lo[p * nchunk + c];
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) {
#pragma unroll
            for (int t = 0; t < NMAT; ++t) {
                const u32x8 w = lds_chunk32(iv.ptr(t, r0 + rr, c * 32), lane);
                const uint32_t sc2 = *reinterpret_cast<const uint16_t*>(iv.ptr(t, r0 + rr, soff + 2 * c));
                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                for (int p = 0; p < 4; ++p) dot8_e2m1(s0, s1, w.v[p], a[p]);
                acc[t][rr] = fmaf(s0 + s1, ue4m3_to_f32(sc2 & 0xffu), acc[t][rr]);
                s0 = 0.f; s1 = 0.f;
#pragma unroll
                for (int p = 4; p < 8; ++p) dot8_e2m1(s0, s1, w.v[p], a[p]);
                acc[t][rr] = fmaf(s0 + s1, ue4m3_to_f32(sc2 >> 8), acc[t][rr]);
            }
        }
    }
#pragma unroll
    for (int t = 0; t < NMAT; ++t)
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) acc[t][rr] = warp_sum(acc[t][rr]);
    if (lane == 0) {
#pragma unroll
        for (int rr = 0; rr < RR; ++rr) {
            if (NMAT == 2) finish_swiglu(ro, n + rr, ro.alpha * (acc[0][rr] * ro.inv), ro.alpha * (acc[NMAT - 1][rr] * ro.inv));
            else finish_row(ro, n + rr, ro.alpha * (acc[0][rr] * ro.inv), res[rr]);
        }
    }
}

// BF16 (lm_head): chunk = 16 elements; acc += dot16 per chunk in increasing order (= gemv_bf16_kernel)
__device__ __forceinline__ float bf16_row(const ItemView& iv, const uint4* a_lo, int nchunk, int r, int lane) {
    float acc = 0.f;
#pragma unroll 2
    for (int c = lane; c < nchunk; c += 32) {
        const u32x8 w = lds_chunk32(iv.ptr(0, r, c * 32), lane);
        float s0 = 0.f, s1 = 0.f;
        dot8_bf16(s0, s1, &w.v[0], a_lo[c]);
        dot8_bf16(s0, s1, &w.v[4], a_lo[nchunk + c]);
        acc += s0 + s1;
    }
    return warp_sum(acc);
}

struct Shared {
    uint64_t full_bar[kPhases][kMaxSlots];
    uint64_t empty_bar[kMaxSlots];
    pgk_step_op_t op[2];             // descriptor of the current / next op (double buffered)
    uint16_t q_s[8][128];
    uint16_t new_s[2][128];
    uint32_t red[80];
    float inv_s[4];
    unsigned long long key;
    int abort_flag;
    float scales[4 * kScalePad];     // FP8 block scales of this CTA's rows: [matrix 0/1][row block 0/1][k block]
    uint16_t xout[kXMax];            // this CTA's partial sums of a row-parallel op (tensor parallel)
    float xin[kMaxWorld][kXMax];     // every rank's partial sums of this CTA's rows
};

// ---------------------------------------------------------------- tensor-parallel exchange over NVLink peer memory
// LL protocol: every 8-byte word carries {payload, exchange number}; an 8-byte store is delivered atomically, so a
// word whose number matches is valid -- no fence, no separate flag, one NVLink hop of latency.
__device__ __forceinline__ void st_ll(unsigned char* p, uint32_t data, uint32_t seq) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(p), "r"(data), "r"(seq) : "memory");
}
__device__ __forceinline__ uint32_t ld_ll_wait(const unsigned char* p, uint32_t seq, Abort ab) {
    uint32_t d, f;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d), "=r"(f) : "l"(p) : "memory");
    if (f == seq) return d;
    const long long t0 = clock64();
    while (true) {
        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d), "=r"(f) : "l"(p) : "memory");
        if (f == seq) return d;
        if (*ab.s_abort) return 0u;
        if (clock64() - t0 > kSpinLimit) { *ab.s_abort = 1; atomicCAS(ab.gerr, 0, 400); return 0u; }
    }
}
// after a row-parallel GEMV whose partial sums of rows [lo, lo + len) sit in sh.xout: push them to every rank, collect
// every rank's partial sums of the same rows, out[n] = bf16(sum over ranks in rank order) (+ residual, one rounding)
__device__ __forceinline__ void xchg_rows(const StepHeader& h, Shared& sh, const pgk_step_op_t& op, int token, int lo, int len,
                                          uint32_t seq, int tid, Abort ab) {
    const int world = h.tp_world;
    const size_t par_off = (size_t)(seq & 1u) * kMaxWorld * h.ll_words;
    consumer_sync();                                   // every row of the CTA is in sh.xout
    for (int idx = tid; idx < len * world; idx += kCT) {
        const int peer = idx / len, j = idx - peer * len;
        st_ll(h.ll[peer] + (par_off + (size_t)h.tp_rank * h.ll_words + lo + j) * 8, sh.xout[j], seq);
    }
    for (int idx = tid; idx < len * world; idx += kCT) {
        const int src = idx / len, j = idx - src * len;
        const uint32_t d = ld_ll_wait(h.ll[h.tp_rank] + (par_off + (size_t)src * h.ll_words + lo + j) * 8, seq, ab);
        sh.xin[src][j] = bf16_bits_to_f32(d & 0xffffu);
    }
    consumer_sync();
    const uint16_t* res = op.residual ? (const uint16_t*)op.residual + ((op.flags & F_RES_EMBED) ? (size_t)token * op.N : 0) : nullptr;
    uint16_t* out = (uint16_t*)op.out;
    for (int j = tid; j < len; j += kCT) {
        float acc = 0.f;
        for (int r = 0; r < world; ++r) acc += sh.xin[r][j];
        out[lo + j] = res ? hadd_bf16_bits(acc, ld_act2(res + lo + j)) : f32_to_bf16_bits(acc);
    }
}

// slot bookkeeping without divisions: (slot index in the ring, ring round)
struct SlotPos {
    int s; uint32_t r;
    __device__ __forceinline__ void advance(int n, int ns) { s += n; while (s >= ns) { s -= ns; ++r; } }
};

// ---------------------------------------------------------------- one GEMV-type op, consumer side
// warp 0 copies the next op's descriptor into the other shared buffer (read after the next grid barrier); called after
// the staging so that its L2 round trip is off the critical path of the op boundary
__device__ __forceinline__ void prefetch_next_op(const StepHeader& h, Shared& sh, int oi, int tid) {
    constexpr int kOpWords = (int)(sizeof(pgk_step_op_t) / 4);
    if (tid < kOpWords && oi + 1 < h.n_ops)
        reinterpret_cast<uint32_t*>(&sh.op[(oi + 1) & 1])[tid] = reinterpret_cast<const uint32_t*>(h.ops + oi + 1)[tid];
}

__device__ __forceinline__ void gemv_op(const pgk_step_op_t& op, const GemvPlan& pl, const StepHeader& h, Shared& sh,
                                        unsigned char* ring, unsigned char* act, SlotPos base, int token, int tid, Abort ab,
                                        float& best_v, int& best_i, unsigned long long* tr, int oi) {
    const int lane = tid & 31, warp = tid >> 5;
    const int K = op.K, N = op.N, kind = op.kind, flags = op.flags;
    const uint16_t* A = (const uint16_t*)op.act + ((flags & F_ACT_EMBED) ? (size_t)token * K : 0);
    const uint16_t* gam = (const uint16_t*)op.gamma;
    const int rb0 = (pl.item0 * pl.R) >> 7;
    ScaleJob sj = {};
    if (kind == OP_FP8) {
        sj.S0 = (const uint16_t*)op.S0; sj.S1 = (const uint16_t*)op.S1;
        sj.SK = (K + 127) >> 7; sj.nrb = (N + 127) >> 7; sj.rb0 = rb0;
        sj.count = ((flags & F_SWIGLU) ? 2 : 1) * 2 * sj.SK;
        sj.dst = sh.scales;
    }
    if (kind == OP_FP8)        stage_act<4, true>(A, K, gam, op.eps, act, sh.red, sh.inv_s, tid, sj, tr);
    else if (kind == OP_NVF4R) stage_act<8, true>(A, K, gam, op.eps, act, sh.red, sh.inv_s, tid, sj, tr);
    else                       stage_act<2, false>(A, K, gam, op.eps, act, sh.red, sh.inv_s, tid, sj, tr);
    if (tr && tid == 0) tr[1] = globaltimer_ns();
    prefetch_next_op(h, sh, oi, tid);
    RowOut ro;
    ro.C = (uint16_t*)op.out;
    ro.residual = op.residual ? (const uint16_t*)op.residual + ((flags & F_RES_EMBED) ? (size_t)token * N : 0) : nullptr;
    ro.inv = (kind == OP_BF16) ? 1.f : sh.inv_s[0];
    ro.alpha = op.alpha;
    ro.flags = flags;
    ro.xout = (flags & F_XCHG) ? sh.xout : nullptr;
    ro.xlo = pl.item0 * pl.R;
    const uint4* a_lo = reinterpret_cast<const uint4*>(act);
    const int nchunk = (kind == OP_FP8) ? (K >> 5) : (kind == OP_NVF4R ? (K >> 6) : (K >> 4));
    ItemView iv;
    iv.ring = ring; iv.ns = h.ns; iv.R = pl.R; iv.RB = pl.RB; iv.nseg = pl.nseg; iv.seg_bytes = pl.seg_bytes; iv.spi = pl.spi;
    SlotPos sp = base;
    sp.advance(warp * pl.spi, h.ns);
    for (int i = warp; i < pl.n_items; i += kCW) {
        iv.qmod = sp.s;
        {
            SlotPos w = sp;
            for (int j = 0; j < pl.spi; ++j) {
                mbar_wait_b(&sh.full_bar[w.r & (kPhases - 1)][w.s], (w.r / kPhases) & 1, ab, 200 + kind);
                w.advance(1, h.ns);
            }
        }
        const int n0 = (pl.item0 + i) * pl.R;
        const int rows = min(pl.R, N - n0);
        if (kind == OP_FP8) {
            if (flags & F_SWIGLU) {
                for (int r = 0; r < rows; ++r) fp8_rows<2, 1>(sh.scales, kScalePad, rb0, iv, a_lo, nchunk, n0 + r, r, lane, ro);
            } else {
                int r = 0;
                for (; r + 2 <= rows; r += 2) fp8_rows<1, 2>(sh.scales, kScalePad, rb0, iv, a_lo, nchunk, n0 + r, r, lane, ro);
                if (r < rows) fp8_rows<1, 1>(sh.scales, kScalePad, rb0, iv, a_lo, nchunk, n0 + r, r, lane, ro);
            }
        } else if (kind == OP_NVF4R) {
            if (flags & F_SWIGLU) {
                for (int r = 0; r < rows; ++r) nvf4_rows<2, 1>(op, iv, a_lo, nchunk, n0 + r, r, lane, ro);
            } else {
                int r = 0;
                for (; r + 2 <= rows; r += 2) nvf4_rows<1, 2>(op, iv, a_lo, nchunk, n0 + r, r, lane, ro);
                if (r < rows) nvf4_rows<1, 1>(op, iv, a_lo, nchunk, n0 + r, r, lane, ro);
            }
        } else {
            for (int r = 0; r < rows; ++r) {
                const uint16_t res = (lane == 0 && ro.residual && !ro.xout) ? ld_act2(ro.residual + n0 + r) : (uint16_t)0;
                const float v = bf16_row(iv, a_lo, nchunk, r, lane);
                if (lane == 0) {
                    const uint16_t ob = finish_row(ro, n0 + r, v, res);
                    if (flags & F_ARGMAX) {
                        const float rv = bf16_bits_to_f32(ob);
                        if (rv > best_v) { best_v = rv; best_i = n0 + r; }   // rows ascend within a warp: first max kept
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) {
            SlotPos w = sp;
            for (int j = 0; j < pl.spi; ++j) { mbar_arrive(&sh.empty_bar[w.s]); w.advance(1, h.ns); }
        }
        sp.advance(kCW * pl.spi, h.ns);
    }
    if (tr && tid == 0) tr[2] = globaltimer_ns();
}

// ---------------------------------------------------------------- attention op, consumer side
// one warp, one unit of <= 16 KV positions, RT query rows: lane owns head dims 4*lane .. 4*lane+3; the 32 score
// partials of an iteration (U = 32/RT positions x RT rows) are reduce-scattered so that lane v owns the score of
// (position v / RT, row v % RT).
template <int RT>
struct AttnState {
    float m_own, l_own;
    float o[RT][4];
};

template <int RT>
__device__ __forceinline__ void attn_unit(AttnState<RT>& st, const unsigned char* tile, int npos, const uint32_t (&q)[RT][2],
                                          float scale_log2, int lane) {
    constexpr int U = 32 / RT;
    const unsigned char* kp = tile + lane * 8;
    const unsigned char* vp = kp + kAttnT * 256;
#pragma unroll 1
    for (int it = 0; it < kAttnT / U; ++it) {
        if (it * U >= npos) break;            // warp-uniform
        uint2 kq[U], vq[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int j = it * U + u;
            if (j < npos) {
                kq[u] = *reinterpret_cast<const uint2*>(kp + j * 256);
                vq[u] = *reinterpret_cast<const uint2*>(vp + j * 256);
            } else {
                kq[u] = make_uint2(0u, 0u); vq[u] = make_uint2(0u, 0u);
            }
        }
        float val[32];
#pragma unroll
        for (int u = 0; u < U; ++u) {
#pragma unroll
            for (int r = 0; r < RT; ++r) {
                float a0 = 0.f, a1 = 0.f;
                fhfma_bf16_lo(a0, q[r][0], kq[u].x); fhfma_bf16_hi(a1, q[r][0], kq[u].x);
                fhfma_bf16_lo(a0, q[r][1], kq[u].y); fhfma_bf16_hi(a1, q[r][1], kq[u].y);
                val[u * RT + r] = a0 + a1;
            }
        }
        float tot;
        {
            const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2, b1 = lane & 1;
            float v16[16], v8[8], v4[4], v2[2];
#pragma unroll
            for (int i = 0; i < 16; ++i)
                v16[i] = (b16 ? val[i + 16] : val[i]) + __shfl_xor_sync(0xffffffffu, b16 ? val[i] : val[i + 16], 16);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                v8[i] = (b8 ? v16[i + 8] : v16[i]) + __shfl_xor_sync(0xffffffffu, b8 ? v16[i] : v16[i + 8], 8);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                v4[i] = (b4 ? v8[i + 4] : v8[i]) + __shfl_xor_sync(0xffffffffu, b4 ? v8[i] : v8[i + 4], 4);
#pragma unroll
            for (int i = 0; i < 2; ++i)
                v2[i] = (b2 ? v4[i + 2] : v4[i]) + __shfl_xor_sync(0xffffffffu, b2 ? v4[i] : v4[i + 2], 2);
            tot = (b1 ? v2[1] : v2[0]) + __shfl_xor_sync(0xffffffffu, b1 ? v2[0] : v2[1], 1);
        }
        const bool own_valid = it * U + lane / RT < npos;
        const float s_own = own_valid ? tot * scale_log2 : kNeg;
        float mx = s_own;
#pragma unroll
        for (int off = RT; off < 32; off <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
        const float mn = fmaxf(st.m_own, mx);
        const float corr_own = exp2f(st.m_own - mn);
        const float p_own = own_valid ? exp2f(s_own - mn) : 0.f;
        float ls = p_own;
#pragma unroll
        for (int off = RT; off < 32; off <<= 1) ls += __shfl_xor_sync(0xffffffffu, ls, off);
        st.l_own = st.l_own * corr_own + ls;
        st.m_own = mn;
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            const float corr = __shfl_sync(0xffffffffu, corr_own, r);
#pragma unroll
            for (int e = 0; e < 4; ++e) st.o[r][e] *= corr;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const float vf[4] = {bf16lo(vq[u].x), bf16hi(vq[u].x), bf16lo(vq[u].y), bf16hi(vq[u].y)};
#pragma unroll
            for (int r = 0; r < RT; ++r) {
                const float pj = __shfl_sync(0xffffffffu, p_own, u * RT + r);
#pragma unroll
                for (int e = 0; e < 4; ++e) st.o[r][e] = fmaf(pj, vf[e], st.o[r][e]);
            }
        }
    }
}

// one warp: [RMSNorm] + NeoX rotation of one 128-wide head row (same arithmetic as norm_rope_head128 in attention.cu)
__device__ __forceinline__ void step_norm_rope128(const uint16_t* row, const uint16_t* gamma, float eps, const float* cs,
                                                  const float* sn, int lane, uint16_t* dst_s, uint16_t* dst_g) {
    float x[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) x[i] = bf16_bits_to_f32(ld_act2(row + lane + 32 * i));
    if (gamma) {
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            float part = warp_sum(fmaf(x[w], x[w], 0.f));
            if (lane == w) tot = part;
        }
        const float inv = rsqrtf(warp_sum(tot) / 128.f + eps);
#pragma unroll
        for (int i = 0; i < 4; ++i) x[i] = round_bf16(x[i] * inv * bf16_bits_to_f32(__ldg(gamma + lane + 32 * i)));
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float c = __ldg(cs + lane + 32 * i), s = __ldg(sn + lane + 32 * i);
        const float x0 = x[i], x1 = x[i + 2];
        const uint16_t y0 = f32_to_bf16_bits(x0 * c - x1 * s), y1 = f32_to_bf16_bits(x1 * c + x0 * s);
        dst_s[lane + 32 * i] = y0; dst_s[lane + 32 * i + 64] = y1;
        if (dst_g) { dst_g[lane + 32 * i] = y0; dst_g[lane + 32 * i + 64] = y1; }
    }
}

template <int RT>
__device__ __forceinline__ void attn_op(const pgk_step_op_t& op, const AttnPlan& ap, const StepHeader& h, Shared& sh,
                                        unsigned char* ring, unsigned char* act, SlotPos base, int cta, int pos, int ctx, int tid,
                                        Abort ab, unsigned long long* tr, int oi) {
    const int lane = tid & 31, warp = tid >> 5;
    const int hq = op.hq, n_kv = op.n_kv, G = hq / n_kv, max_seq = op.max_seq;
    const uint16_t* qkv = (const uint16_t*)op.act;
    uint16_t* kc = (uint16_t*)op.k_cache;
    uint16_t* vc = (uint16_t*)op.v_cache;
    const bool has_new = pos < ctx;
    const int tpos = min(pos, max_seq - 1);
    const float* cs = op.cos_table + (size_t)tpos * 128;
    const float* sn = op.sin_table + (size_t)tpos * 128;
    const float scale = op.scale > 0.f ? op.scale : 0.08838834764831845f;   // 1/sqrt(128)
    const float scale_log2 = scale * 1.4426950408889634f;
    float* part = reinterpret_cast<float*>(act);     // [8][RT][kPartStride]
    const int kvh = ap.kvh, u0 = ap.u0, u1 = ap.u1;
    if (u0 >= u1) { prefetch_next_op(h, sh, oi, tid); return; }   // CTA-uniform: more CTAs than tiles in this head's group
    // ---- the group's query rows and the new K / V row
    const int tw = has_new ? pos / kAttnT : 0;
    const bool writer = pos < max_seq && tw >= u0 && tw < u1;
    if (warp < RT) {
        if (warp < G) step_norm_rope128(qkv + (size_t)(kvh * G + warp) * 128, (const uint16_t*)op.q_gamma, op.attn_eps, cs, sn,
                                        lane, sh.q_s[warp], nullptr);
        else for (int d = lane; d < 128; d += 32) sh.q_s[warp][d] = 0;
    } else if (warp == RT) {
        uint16_t* dst = writer ? kc + ((size_t)kvh * max_seq + pos) * 128 : nullptr;
        step_norm_rope128(qkv + (size_t)(hq + kvh) * 128, (const uint16_t*)op.k_gamma, op.attn_eps, cs, sn, lane,
                          sh.new_s[0], dst);
    } else if (warp == RT + 1) {
        const uint16_t* src = qkv + (size_t)(hq + n_kv + kvh) * 128;
        uint16_t* dst = writer ? vc + ((size_t)kvh * max_seq + pos) * 128 : nullptr;
        for (int d = lane; d < 128; d += 32) { const uint16_t v = ld_act2(src + d); sh.new_s[1][d] = v; if (dst) dst[d] = v; }
    }
    consumer_sync();
    if (tr && tid == 0) tr[1] = globaltimer_ns();
    prefetch_next_op(h, sh, oi, tid);
    uint32_t q[RT][2];
#pragma unroll
    for (int r = 0; r < RT; ++r) {
        const uint2 v = *reinterpret_cast<const uint2*>(&sh.q_s[r][lane * 4]);
        q[r][0] = v.x; q[r][1] = v.y;
    }
    AttnState<RT> st;
    st.m_own = kNeg; st.l_own = 0.f;
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int e = 0; e < 4; ++e) st.o[r][e] = 0.f;
    SlotPos sp = base;
    sp.advance(warp, h.ns);
    for (int t = u0 + warp; t < u1; t += kCW) {
        mbar_wait_b(&sh.full_bar[sp.r & (kPhases - 1)][sp.s], (sp.r / kPhases) & 1, ab, 204);
        unsigned char* tile = ring + (size_t)sp.s * kSlotB;
        const int t0 = t * kAttnT;
        const int npos = min(kAttnT, ctx - t0);
        const bool patch = has_new && pos >= t0 && pos < t0 + npos;     // warp-uniform
        if (patch) {
            *reinterpret_cast<uint2*>(tile + (size_t)(pos - t0) * 256 + lane * 8) =
                *reinterpret_cast<const uint2*>(&sh.new_s[0][lane * 4]);
            *reinterpret_cast<uint2*>(tile + kAttnT * 256 + (size_t)(pos - t0) * 256 + lane * 8) =
                *reinterpret_cast<const uint2*>(&sh.new_s[1][lane * 4]);
            __syncwarp();
        }
        attn_unit<RT>(st, tile, npos, q, scale_log2, lane);
        if (patch) fence_proxy_async();   // the patch was a generic-proxy write; the slot's next fill is a TMA write
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty_bar[sp.s]);
        sp.advance(kCW, h.ns);
    }
    if (tr && tid == 0) tr[2] = globaltimer_ns();
    // ---- fold the warps' states (8 warps at a time through shared memory), write the CTA's record
    constexpr int NX = (RT * 128 + kCT - 1) / kCT;       // outputs (row, dim) per thread
    float Mr[NX], Or[NX], Lr[NX];
#pragma unroll
    for (int x = 0; x < NX; ++x) { Mr[x] = kNeg; Or[x] = 0.f; Lr[x] = 0.f; }
    for (int grp = 0; grp < (kCW + 7) / 8; ++grp) {
        const int nw = min(8, kCW - grp * 8);
        if ((warp >> 3) == grp) {
            const int w8 = warp & 7;
#pragma unroll
            for (int r = 0; r < RT; ++r)
                *reinterpret_cast<float4*>(&part[(w8 * RT + r) * kPartStride + 4 * lane]) =
                    make_float4(st.o[r][0], st.o[r][1], st.o[r][2], st.o[r][3]);
            if (lane < RT) {
                part[(w8 * RT + lane) * kPartStride + 128] = st.m_own;
                part[(w8 * RT + lane) * kPartStride + 129] = st.l_own;
            }
        }
        consumer_sync();
#pragma unroll
        for (int x = 0; x < NX; ++x) {
            const int idx = tid + x * kCT;
            if (idx < RT * 128) {
                const int r = idx >> 7, d = idx & 127;
                float mn = kNeg;
                for (int w = 0; w < nw; ++w) mn = fmaxf(mn, part[(w * RT + r) * kPartStride + 128]);
                float ov = 0.f, lv = 0.f;
                for (int w = 0; w < nw; ++w) {
                    const float c = exp2f(part[(w * RT + r) * kPartStride + 128] - mn);
                    ov = fmaf(part[(w * RT + r) * kPartStride + d], c, ov);
                    lv = fmaf(part[(w * RT + r) * kPartStride + 129], c, lv);
                }
                const float Mn = fmaxf(Mr[x], mn);
                const float c1 = exp2f(Mr[x] - Mn), c2 = exp2f(mn - Mn);
                Or[x] = Or[x] * c1 + ov * c2;
                Lr[x] = Lr[x] * c1 + lv * c2;
                Mr[x] = Mn;
            }
        }
        consumer_sync();
    }
    float* rec = h.attn_ws + (size_t)cta * RT * kRec;
#pragma unroll
    for (int x = 0; x < NX; ++x) {
        const int idx = tid + x * kCT;
        if (idx < RT * 128) {
            const int r = idx >> 7, d = idx & 127;
            __stcg(rec + r * kRec + d, Or[x]);
            if (d == 0) { __stcg(rec + r * kRec + 128, Mr[x]); __stcg(rec + r * kRec + 129, Lr[x]); }
        }
    }
    if (tr && tid == 0) tr[3] = globaltimer_ns();
}

// after the grid barrier: out[head, :] = fold over the CTAs of the head's KV group, in CTA order
template <int RT>
__device__ __forceinline__ void attn_combine(const pgk_step_op_t& op, const StepHeader& h, int cta, int ncta, int ctx, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    const int hq = op.hq, n_kv = op.n_kv, G = hq / n_kv;
    const int ntile = (ctx + kAttnT - 1) / kAttnT;
    const int items = hq * 4;
    const int i0 = (int)part_lo((unsigned)items, (unsigned)cta, (unsigned)ncta);
    const int i1 = (int)part_lo((unsigned)items, (unsigned)cta + 1u, (unsigned)ncta);
    uint16_t* out = (uint16_t*)op.out;
    for (int i = i0 + warp; i < i1; i += kCW) {
        const int head = i >> 2, qd = i & 3;
        const int kvh = head / G, g = head - kvh * G;
        const int cf = attn_cfirst(kvh, n_kv, ncta);
        const int gs = min(attn_cfirst(kvh + 1, n_kv, ncta) - cf, kAttnMaxGroup);
        const bool all_live = ntile >= gs;    // with fewer tiles than CTAs only some CTAs of the group hold a record
        const float* base = h.attn_ws + (size_t)cf * RT * kRec + (size_t)g * kRec;
        // one pass, records in CTA order, 16 records' loads in flight at a time; online rescale of (O, L)
        float M = kNeg, O = 0.f, L = 0.f;
        for (int l0 = 0; l0 < gs; l0 += 16) {
            float mv[16], lv[16], ov[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const int lc = l0 + u;
                const bool has = lc < gs && (all_live || (lc * ntile / gs < (lc + 1) * ntile / gs));
                const float* rr = base + (size_t)lc * RT * kRec;
                mv[u] = has ? __ldcg(rr + 128) : kNeg;
                lv[u] = has ? __ldcg(rr + 129) : 0.f;
                ov[u] = has ? __ldcg(rr + qd * 32 + lane) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const float Mn = fmaxf(M, mv[u]);
                const float c1 = exp2f(M - Mn), c2 = exp2f(mv[u] - Mn);
                O = O * c1 + ov[u] * c2;
                L = L * c1 + lv[u] * c2;
                M = Mn;
            }
        }
        out[(size_t)head * 128 + qd * 32 + lane] = f32_to_bf16_bits(L > 0.f ? O / L : 0.f);
    }
}

// ---------------------------------------------------------------- producers: one thread per producer warp; producer p
// fills the slots whose sequence number is p (mod kNP).  Both walk the whole op list (same plans as the consumers).
__device__ __forceinline__ void producer(const StepHeader& h, Shared& sh, unsigned char* ring, int p, int cta, int ncta,
                                         int ctx_in, Abort ab) {
    const uint64_t pol = make_evict_first_policy();
    const uint32_t ring_u = smem_u32(ring);
    const int ns = h.ns;
    SlotPos sp = {0, 0u};
    int turn = 0;                      // sequence number of the next slot, mod kNP
    // Outstanding fills are capped: every byte in flight queues in front of the loads on the op boundaries' critical
    // path (grid-barrier polls, activation vectors) -- bandwidth needs ~35 KB in flight per SM, not the whole ring.
    uint32_t win_bar[3] = {0u, 0u, 0u}, win_par[3] = {0u, 0u, 0u};     // full barrier + parity of the last fills, newest first
    const int depth = h.inflight < 1 ? 1 : (h.inflight > 3 ? 3 : h.inflight);
    // claim the next slot if it is this producer's: returns its shared address (0 = not mine) and arms the barrier
    auto next_slot = [&](uint32_t bytes, uint32_t& bar) -> uint32_t {
        uint32_t dst = 0;
        if (turn == p) {
            const uint32_t ob = depth == 1 ? win_bar[0] : (depth == 2 ? win_bar[1] : win_bar[2]);
            const uint32_t op_ = depth == 1 ? win_par[0] : (depth == 2 ? win_par[1] : win_par[2]);
            if (ob) mbar_wait_b(reinterpret_cast<uint64_t*>(__cvta_shared_to_generic(ob)), op_, ab, 301);
            mbar_wait_b(&sh.empty_bar[sp.s], (sp.r & 1) ^ 1, ab, 300);
            bar = smem_u32(&sh.full_bar[sp.r & (kPhases - 1)][sp.s]);
            mbar_expect_tx_u32(bar, bytes);
            dst = ring_u + (uint32_t)sp.s * kSlotB;
            win_bar[2] = win_bar[1]; win_par[2] = win_par[1];
            win_bar[1] = win_bar[0]; win_par[1] = win_par[0];
            win_bar[0] = bar; win_par[0] = (sp.r / kPhases) & 1;
        }
        turn = (turn + 1 == kNP) ? 0 : turn + 1;
        sp.advance(1, ns);
        return dst;
    };
    for (int oi = 0; oi < h.n_ops; ++oi) {
        const pgk_step_op_t* op = h.ops + oi;
        const int kind = op->kind;
        if (kind == OP_ATTN) {
            const int max_seq = op->max_seq;
            const int ctx = attn_ctx(ctx_in, max_seq);
            const AttnPlan a = attn_plan(op->n_kv, ctx, cta, ncta);
            const unsigned char* kc = (const unsigned char*)op->k_cache + (size_t)a.kvh * max_seq * 256;
            const unsigned char* vc = (const unsigned char*)op->v_cache + (size_t)a.kvh * max_seq * 256;
            for (int t = a.u0; t < a.u1; ++t) {
                const int t0 = t * kAttnT;
                const uint32_t bytes = (uint32_t)min(kAttnT, ctx - t0) * 256u;
                uint32_t bar = 0;
                const uint32_t dst = next_slot(2 * bytes, bar);
                if (dst) {
                    bulk_copy_plain(dst, kc + (size_t)t0 * 256, bytes, bar);
                    bulk_copy_plain(dst + kAttnT * 256, vc + (size_t)t0 * 256, bytes, bar);
                }
            }
            continue;
        }
        const int N = op->N;
        const GemvPlan pl = gemv_plan(kind, op->flags, op->K, N, cta, ncta);
        const unsigned char* W0 = (const unsigned char*)op->W0;
        const unsigned char* W1 = (const unsigned char*)op->W1;
        if (pl.spi == 1) {
            const uint32_t item_bytes = (uint32_t)pl.R * pl.RB;
            const unsigned char* s0 = W0 + (size_t)pl.item0 * item_bytes;
            const unsigned char* s1 = W1 + (size_t)pl.item0 * item_bytes;
            const int last = pl.items_total - 1 - pl.item0;          // local index of the matrix's last (possibly short) item
            const uint32_t last_bytes = (uint32_t)(N - (pl.items_total - 1) * pl.R) * pl.RB;
            for (int i = 0; i < pl.n_items; ++i) {
                const uint32_t bytes = (i == last) ? last_bytes : item_bytes;
                uint32_t bar = 0;
                const uint32_t dst = next_slot(bytes * pl.nmat, bar);
                if (dst) {
                    bulk_copy_hint(dst, s0, bytes, bar, pol);
                    if (pl.nmat == 2) bulk_copy_hint(dst + item_bytes, s1, bytes, bar, pol);
                }
                s0 += item_bytes; s1 += item_bytes;
            }
        } else {
            for (int i = 0; i < pl.n_items; ++i) {
                const size_t row = (size_t)(pl.item0 + i) * pl.RB;
                for (int t = 0; t < pl.nmat; ++t)
                    for (int sgi = 0; sgi < pl.nseg; ++sgi) {
                        uint32_t bar = 0;
                        const uint32_t dst = next_slot((uint32_t)pl.seg_bytes, bar);
                        if (dst) bulk_copy_hint(dst, (t ? W1 : W0) + row + (size_t)sgi * pl.seg_bytes, (uint32_t)pl.seg_bytes, bar, pol);
                    }
            }
        }
    }
}

__global__ void __launch_bounds__(kStepThreads, 1) step_kernel(const StepHeader h) {
    extern __shared__ __align__(1024) unsigned char dyn[];
    __shared__ Shared sh;
    unsigned char* ring = dyn;
    unsigned char* act = dyn + (size_t)h.ns * kSlotB;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cta = blockIdx.x, ncta = gridDim.x;
    griddep_wait();
    for (int i = tid; i < h.ns * (kPhases + 1); i += kStepThreads) {
        const int slot = i % h.ns, ph = i / h.ns;
        mbar_init(ph == kPhases ? &sh.empty_bar[slot] : &sh.full_bar[ph][slot], 1);
    }
    constexpr int kOpWords = (int)(sizeof(pgk_step_op_t) / 4);
    if (tid < kOpWords) reinterpret_cast<uint32_t*>(&sh.op[0])[tid] = reinterpret_cast<const uint32_t*>(h.ops)[tid];
    if (tid == 0) { sh.key = 0ull; sh.abort_flag = 0; }
    mbar_fence_init();
    __syncthreads();
    Abort ab = {&sh.abort_flag, h.error};
    const int pos = *h.pos_buf;
    const int ctx_raw = *h.ctx_buf;
    int token = *h.token_buf;
    if (token < 0 || token >= h.vocab) token = 0;
    const unsigned parity = h.bar[2] & 1u;
    const uint32_t ll_base = h.tp_world > 1 ? *h.ll_epoch : 0u;      // exchanges completed before this step

    if (warp >= kCW) {                                   // ---- producer warps
        if (lane == 0) producer(h, sh, ring, warp - kCW, cta, ncta, ctx_raw, ab);   // attention ops clamp the context themselves
        return;
    }
    // ---- consumers
    unsigned* ctr = h.bar + parity;
    if (cta == 0 && tid == 0) h.bar[parity ^ 1u] = 0u;   // the other parity's counter: idle during this step
    unsigned target = 0;
    SlotPos base = {0, 0u};
    float best_v = -INFINITY;
    int best_i = 0x7fffffff;
    bool any_argmax = false;
    uint32_t xi = 0;                                     // exchanges done in this step
    for (int oi = 0; oi < h.n_ops; ++oi) {
        unsigned long long* tr = (cta == 0) ? h.trace + (size_t)oi * kTracePts : nullptr;
        if (oi > 0) { target += (unsigned)ncta; grid_sync(ctr, target, tid, ab, tr ? tr - kTracePts + 4 : nullptr, h.bar_mode); }
        if (tr && tid == 0) tr[0] = globaltimer_ns();
        const pgk_step_op_t& op = sh.op[oi & 1];
        if (op.kind == OP_ATTN) {
            const int ctx = attn_ctx(ctx_raw, op.max_seq);
            const AttnPlan ap = attn_plan(op.n_kv, ctx, cta, ncta);
            const int G = op.hq / op.n_kv;
            if (G <= 4) attn_op<4>(op, ap, h, sh, ring, act, base, cta, pos, ctx, tid, ab, tr, oi);
            else        attn_op<8>(op, ap, h, sh, ring, act, base, cta, pos, ctx, tid, ab, tr, oi);
            base.advance(ap.u1 - ap.u0, h.ns);
            target += (unsigned)ncta;
            grid_sync(ctr, target, tid, ab, tr ? tr + 5 : nullptr, h.bar_mode);
            if (tr && tid == 0) tr[6] = globaltimer_ns();
            if (G <= 4) attn_combine<4>(op, h, cta, ncta, ctx, tid);
            else        attn_combine<8>(op, h, cta, ncta, ctx, tid);
        } else {
            const GemvPlan pl = gemv_plan(op.kind, op.flags, op.K, op.N, cta, ncta);
            gemv_op(op, pl, h, sh, ring, act, base, token, tid, ab, best_v, best_i, tr, oi);
            if (op.flags & F_XCHG) {
                const int lo = pl.item0 * pl.R;
                xchg_rows(h, sh, op, token, lo, min(op.N, (pl.item0 + pl.n_items) * pl.R) - lo, ll_base + (++xi), tid, ab);
            }
            base.advance(pl.n_items * pl.spi, h.ns);
            any_argmax |= (op.flags & F_ARGMAX) != 0;
        }
    }
    if (any_argmax) {
        if (lane == 0 && best_i != 0x7fffffff) atomicMax(&sh.key, argmax_key(best_v, best_i + h.vocab_offset));
        consumer_sync();
        if (tid == 0 && sh.key != 0ull) atomicMax(h.am_key, sh.key);
    }
    target += (unsigned)ncta;
    unsigned long long* trl = (cta == 0) ? h.trace + (size_t)h.n_ops * kTracePts : nullptr;
    grid_sync(ctr, target, tid, ab, trl ? trl - kTracePts + 4 : nullptr, h.bar_mode);
    if (cta == 0 && tid == 0) {
        trl[0] = globaltimer_ns();
        if (any_argmax) {
            unsigned long long k = atomicExch(h.am_key, 0ull);
            if (h.tp_world > 1) {
                // every rank publishes its best (value, global index) key as two LL words and takes the maximum of all:
                // identical on every rank, lowest index wins ties
                const uint32_t seq = ll_base + (++xi);
                const size_t par_off = (size_t)(seq & 1u) * kMaxWorld * h.ll_words;
                for (int r = 0; r < h.tp_world; ++r) {
                    unsigned char* dst = h.ll[r] + (par_off + (size_t)h.tp_rank * h.ll_words) * 8;
                    st_ll(dst, (uint32_t)(k & 0xffffffffull), seq);
                    st_ll(dst + 8, (uint32_t)(k >> 32), seq);
                }
                unsigned long long best = 0ull;
                for (int r = 0; r < h.tp_world; ++r) {
                    const unsigned char* src = h.ll[h.tp_rank] + (par_off + (size_t)r * h.ll_words) * 8;
                    const unsigned long long lo32 = ld_ll_wait(src, seq, ab), hi32 = ld_ll_wait(src + 8, seq, ab);
                    const unsigned long long kr = (hi32 << 32) | lo32;
                    best = kr > best ? kr : best;
                }
                k = best;
            }
            int tok = 0x7fffffff - (int)(uint32_t)(k & 0xffffffffull);
            if (tok < 0 || tok >= h.vocab || *ab.s_abort) tok = 0;
            *h.next_token = tok;
            *h.token_buf = tok;
        }
        if (h.advance) { *h.pos_buf = pos + 1; *h.ctx_buf = ctx_raw + 1; }
        if (h.tp_world > 1) *h.ll_epoch = ll_base + (uint32_t)h.n_xchg;
        h.bar[2] = h.bar[2] + 1u;
    }
    griddep_launch();
}

// reference NVF4 layout -> row-packed: one thread per (row n, byte j of the packed row); the source column of output
// row n is (n / group) * stride + n % group + col0 (plain: group = N; a gate/up tensor interleaved in groups of 8: 8, 16)
__global__ void nvf4_pack_rows_kernel(const uint8_t* __restrict__ data, const uint8_t* __restrict__ scale,
                                      uint8_t* __restrict__ packed, int K, int N, int ld, int col0, int group, int stride) {
    griddep_wait();
    griddep_launch();
    const int RB = K / 2 + K / 32;
    const size_t total = (size_t)N * RB;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int n = (int)(i / RB), j = (int)(i % RB);
        const size_t col = (size_t)(n / group) * stride + (n % group) + col0;
        packed[i] = j < K / 2 ? data[(size_t)j * ld + col] : scale[(size_t)(j - K / 2) * ld + col];
    }
}
#endif  // __CUDACC__

struct StepProgram {
    StepHeader h;
    pgk_step_op_t* d_ops = nullptr;
    unsigned* d_bar = nullptr;
    float* d_ws = nullptr;
    unsigned long long* d_key = nullptr;
    int* d_err = nullptr;
    unsigned long long* d_trace = nullptr;
    int ctas = 0;
    size_t smem = 0;
};

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_step_create(const pgk_step_op_t* host_ops, int n_ops, int32_t* token_buf, int32_t* pos_buf, int32_t* ctx_buf,
                    int32_t* next_token, int vocab, int advance, void** handle) {
    PGK_REQUIRE(host_ops && n_ops >= 1 && handle, "step_create failed: empty program");
    PGK_REQUIRE(token_buf && pos_buf && ctx_buf && next_token, "step_create failed: decode state buffers must not be NULL");
    const int ctas = sm_count();
    // ---- shared-memory budget: activation area (largest staged vector, or the attention fold scratch), ring = the rest
    size_t act = 1024, ws_floats = 0;
    for (int i = 0; i < n_ops; ++i) {
        const pgk_step_op_t& op = host_ops[i];
        if (op.kind == OP_ATTN) {
            PGK_REQUIRE(op.hq > 0 && op.n_kv > 0 && op.hq % op.n_kv == 0 && op.hq / op.n_kv <= 8,
                        "step_create failed: op %d: attention needs hq %% n_kv == 0 and at most 8 query heads per KV head (hq=%d n_kv=%d)",
                        i, op.hq, op.n_kv);
            PGK_REQUIRE(op.act && op.out && op.cos_table && op.sin_table && op.k_cache && op.v_cache && op.max_seq >= 1,
                        "step_create failed: op %d: attention pointers must not be NULL", i);
            PGK_REQUIRE(op.n_kv <= ctas, "step_create failed: op %d: more KV heads (%d) than CTAs (%d)", i, op.n_kv, ctas);
            const int RT = op.hq / op.n_kv <= 4 ? 4 : 8;
            act = std::max(act, (size_t)8 * RT * kPartStride * sizeof(float));
            ws_floats = std::max(ws_floats, (size_t)ctas * RT * kRec);
            continue;
        }
        PGK_REQUIRE(op.kind == OP_FP8 || op.kind == OP_BF16 || op.kind == OP_NVF4R, "step_create failed: op %d: unknown kind %d", i, op.kind);
        PGK_REQUIRE(op.K >= 64 && op.N >= 1 && op.act && op.W0, "step_create failed: op %d: invalid GEMV (K=%d N=%d)", i, op.K, op.N);
        const int kmul = op.kind == OP_NVF4R ? 512 : (op.kind == OP_FP8 ? 32 : 16);
        PGK_REQUIRE(op.K % kmul == 0, "step_create failed: op %d: K=%d must be a multiple of %d for this weight format", i, op.K, kmul);
        PGK_REQUIRE(op.K / 8 <= 8 * kCT, "step_create failed: op %d: K=%d too large", i, op.K);
        PGK_REQUIRE(op.kind == OP_NVF4R || (op.kind == OP_FP8 ? op.S0 != nullptr : true), "step_create failed: op %d: missing scales", i);
        PGK_REQUIRE(!(op.flags & F_SWIGLU) || (op.kind != OP_BF16 && op.W1 && (op.kind == OP_NVF4R || op.S1) && op.out && !op.residual),
                    "step_create failed: op %d: SWIGLU needs W1 (+S1), out, no residual, FP8 or NVF4 weights", i);
        PGK_REQUIRE(!(op.flags & F_ARGMAX) || op.kind == OP_BF16, "step_create failed: op %d: ARGMAX rides on a BF16 GEMV", i);
        PGK_REQUIRE(op.out || (op.flags & F_ARGMAX), "step_create failed: op %d: out is NULL", i);
        const GemvPlan pl = gemv_plan(op.kind, op.flags, op.K, op.N, 0, ctas);
        PGK_REQUIRE(pl.RB % 16 == 0, "step_create failed: op %d: row bytes %d not a multiple of 16", i, pl.RB);
        PGK_REQUIRE(pl.nseg <= 4 && (pl.nseg == 1 || (pl.RB % pl.nseg == 0 && pl.seg_bytes % 32 == 0)),
                    "step_create failed: op %d: a row of %d bytes cannot be cut into <= 4 equal 32-byte-aligned slot segments", i, pl.RB);
        for (const void* w : {op.W0, (op.flags & F_SWIGLU) ? op.W1 : op.W0})
            PGK_REQUIRE(((uintptr_t)w & 15) == 0, "step_create failed: op %d: weights must be 16-byte aligned", i);
        PGK_REQUIRE(((uintptr_t)op.act & 15) == 0 && (!op.gamma || ((uintptr_t)op.gamma & 15) == 0),
                    "step_create failed: op %d: act / gamma must be 16-byte aligned", i);
        if (op.kind == OP_FP8) {
            const GemvPlan plf = gemv_plan(op.kind, op.flags, op.K, op.N, 0, ctas);
            PGK_REQUIRE((plf.items_total / ctas + 1) * plf.R <= 128 && (op.K + 127) / 128 <= kScalePad &&
                        ((op.flags & F_SWIGLU) ? 2 : 1) * 2 * ((op.K + 127) / 128) <= 2 * kCT,
                        "step_create failed: op %d: more than 128 FP8 rows per CTA or K > 32768 (block scales are staged per CTA)", i);
            // codes 0x7F / 0xFF are +-480 in the reference's tables but NaN for the hardware conversion this kernel uses
            // (common.cuh): such tensors stay on the per-op path, which serves them with the exact table kernel
            bool clean = true;
            if (int rc = fp8_codes_clean(op.W0, (size_t)op.N * op.K, nullptr, &clean)) return rc;
            if (clean && (op.flags & F_SWIGLU)) { if (int rc = fp8_codes_clean(op.W1, (size_t)op.N * op.K, nullptr, &clean)) return rc; }
            PGK_REQUIRE(clean, "step_create failed: op %d: FP8 weights hold E4M3 codes 0x7F/0xFF (table semantics need the per-op path)", i);
        }
        act = std::max(act, (size_t)op.K * 2);
    }
    // ---- tensor parallelism: row-parallel ops exchange their partial sums inside the kernel (LL words over NVLink)
    int n_xchg = 0, tp_rank = 0, tp_world = 1;
    unsigned char* ll[kMaxWorld] = {};
    unsigned* ll_epoch = nullptr;
    size_t ll_words = 0;
    bool any_argmax = false;
    for (int i = 0; i < n_ops; ++i) {
        const pgk_step_op_t& op = host_ops[i];
        any_argmax |= op.kind != OP_ATTN && (op.flags & F_ARGMAX);
        if (op.kind == OP_ATTN || !(op.flags & F_XCHG)) continue;
        if (n_xchg == 0) {
            PGK_REQUIRE(p2p_ll_info(&tp_rank, &tp_world, ll, &ll_epoch, &ll_words) && tp_world > 1,
                        "step_create failed: op %d exchanges partial sums but the peer-memory buffers are not open (pgk_p2p_create / pgk_p2p_open)", i);
        }
        ++n_xchg;
        PGK_REQUIRE(!(op.flags & (F_SWIGLU | F_ARGMAX)) && op.out, "step_create failed: op %d: an exchanged op is a plain row-parallel GEMV", i);
        PGK_REQUIRE((size_t)op.N <= ll_words, "step_create failed: op %d: N=%d exceeds the exchange region (%zu rows)", i, op.N, ll_words);
        const GemvPlan pl = gemv_plan(op.kind, op.flags, op.K, op.N, 0, ctas);
        PGK_REQUIRE((pl.items_total / ctas + 1) * pl.R <= kXMax, "step_create failed: op %d: %d rows per CTA exceed the exchange buffer (%d)",
                    i, (pl.items_total / ctas + 1) * pl.R, kXMax);
    }
    if (n_xchg > 0 && any_argmax) ++n_xchg;      // the (value, index) exchange of the sharded lm_head
    act = (act + 1023) & ~(size_t)1023;
    cudaDeviceProp prop;
    int dev = 0;
    PGK_CUDA(cudaGetDevice(&dev));
    PGK_CUDA(cudaGetDeviceProperties(&prop, dev));
    cudaFuncAttributes fa;
    PGK_CUDA(cudaFuncGetAttributes(&fa, step_kernel));
    const size_t avail = (size_t)prop.sharedMemPerBlockOptin - fa.sharedSizeBytes - 1024;   // 1024: alignment slack of dyn[]
    PGK_REQUIRE(avail > act + 4 * (size_t)kSlotB, "step_create failed: activation area of %zu bytes leaves no room for the ring", act);
    int ns = (int)((avail - act) / kSlotB);
    if (ns > kMaxSlots) ns = kMaxSlots;
    for (int i = 0; i < n_ops; ++i) {
        if (host_ops[i].kind == OP_ATTN) continue;
        const GemvPlan pl = gemv_plan(host_ops[i].kind, host_ops[i].flags, host_ops[i].K, host_ops[i].N, 0, ctas);
        PGK_REQUIRE(pl.spi <= ns && kPhases * ns >= (kCW - 1) * pl.spi + 1,
                    "step_create failed: op %d needs %d ring slots per item, %d available", i, pl.spi, ns);
    }
    StepProgram* sp = new StepProgram();
    sp->ctas = ctas;
    sp->smem = (size_t)ns * kSlotB + act;
    PGK_CUDA(cudaFuncSetAttribute(step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp->smem));
    PGK_CUDA(cudaMalloc(&sp->d_ops, sizeof(pgk_step_op_t) * n_ops));
    PGK_CUDA(cudaMemcpy(sp->d_ops, host_ops, sizeof(pgk_step_op_t) * n_ops, cudaMemcpyHostToDevice));
    PGK_CUDA(cudaMalloc(&sp->d_bar, 4 * sizeof(unsigned)));
    PGK_CUDA(cudaMemset(sp->d_bar, 0, 4 * sizeof(unsigned)));
    PGK_CUDA(cudaMalloc(&sp->d_ws, std::max(ws_floats, (size_t)1) * sizeof(float)));
    PGK_CUDA(cudaMalloc(&sp->d_key, sizeof(unsigned long long)));
    PGK_CUDA(cudaMemset(sp->d_key, 0, sizeof(unsigned long long)));
    PGK_CUDA(cudaMalloc(&sp->d_trace, sizeof(unsigned long long) * (n_ops + 1) * kTracePts));
    PGK_CUDA(cudaMemset(sp->d_trace, 0, sizeof(unsigned long long) * (n_ops + 1) * kTracePts));
    PGK_CUDA(cudaMalloc(&sp->d_err, sizeof(int)));
    PGK_CUDA(cudaMemset(sp->d_err, 0, sizeof(int)));
    PGK_CUDA(cudaDeviceSynchronize());
    StepHeader& h = sp->h;
    {
        const char* e = getenv("PGK_STEP_INFLIGHT");
        h.inflight = e ? atoi(e) : 2;
        const char* b = getenv("PGK_STEP_BAR");
        h.bar_mode = b ? atoi(b) : 0;
    }
    h.ops = sp->d_ops; h.n_ops = n_ops; h.ns = ns; h.act_bytes = (int)act; h.vocab = vocab; h.advance = advance;
    h.token_buf = token_buf; h.pos_buf = pos_buf; h.ctx_buf = ctx_buf; h.next_token = next_token;
    h.bar = sp->d_bar; h.attn_ws = sp->d_ws; h.am_key = sp->d_key; h.error = sp->d_err; h.trace = sp->d_trace;
    h.tp_rank = tp_rank; h.tp_world = n_xchg > 0 ? tp_world : 1; h.vocab_offset = 0; h.n_xchg = n_xchg;
    for (int r = 0; r < kMaxWorld; ++r) h.ll[r] = ll[r];
    h.ll_epoch = ll_epoch; h.ll_words = (unsigned)ll_words;
    *handle = sp;
    return 0;
}

/* tensor parallel: global index of this rank's first lm_head row (the argmax of a sharded lm_head is exchanged as
 * (value, global index) pairs; `vocab` of pgk_step_create is then the GLOBAL vocabulary size) */
int pgk_step_set_vocab_offset(void* handle, int vocab_offset) {
    PGK_REQUIRE(handle && vocab_offset >= 0, "step_set_vocab_offset failed: bad argument");
    ((StepProgram*)handle)->h.vocab_offset = vocab_offset;
    return 0;
}

int pgk_step_run(void* handle, void* stream) {
    PGK_REQUIRE(handle, "step_run failed: NULL program");
    StepProgram* sp = (StepProgram*)handle;
    PGK_CUDA(launch(step_kernel, dim3(sp->ctas), dim3(kStepThreads), sp->smem, (cudaStream_t)stream, sp->h));
    return 0;
}

int pgk_step_error(void* handle, int* host_code) {
    PGK_REQUIRE(handle && host_code, "step_error failed: NULL argument");
    StepProgram* sp = (StepProgram*)handle;
    PGK_CUDA(cudaDeviceSynchronize());
    PGK_CUDA(cudaMemcpy(host_code, sp->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}

int pgk_step_info(void* handle, int* ring_slots, int* act_bytes, int* smem_bytes, int* ctas) {
    PGK_REQUIRE(handle, "step_info failed: NULL program");
    StepProgram* sp = (StepProgram*)handle;
    if (ring_slots) *ring_slots = sp->h.ns;
    if (act_bytes) *act_bytes = sp->h.act_bytes;
    if (smem_bytes) *smem_bytes = (int)sp->smem;
    if (ctas) *ctas = sp->ctas;
    return 0;
}

/* profiling aid: (n_ops + 1) x 8 globaltimer (ns) stamps of CTA 0 for the LAST run.  Per op: [0] start (grid barrier
 * passed), [1] activations staged / attention prologue done, [2] warp 0 finished its items, [3] attention fold done,
 * [4] CTA 0 arrived at the barrier that ends the op, [5] arrival at the attention's inner barrier, [6] attention combine
 * start; row n_ops: [0] = end of the step. */
int pgk_step_trace(void* handle, unsigned long long* host_out, int count) {
    PGK_REQUIRE(handle && host_out, "step_trace failed: NULL argument");
    StepProgram* sp = (StepProgram*)handle;
    PGK_REQUIRE(count >= 1 && count <= (sp->h.n_ops + 1) * kTracePts, "step_trace failed: count %d out of range", count);
    PGK_CUDA(cudaDeviceSynchronize());
    PGK_CUDA(cudaMemcpy(host_out, sp->d_trace, sizeof(unsigned long long) * count, cudaMemcpyDeviceToHost));
    return 0;
}

int pgk_step_destroy(void* handle) {
    if (!handle) return 0;
    StepProgram* sp = (StepProgram*)handle;
    cudaFree(sp->d_trace);
    cudaFree(sp->d_ops); cudaFree(sp->d_bar); cudaFree(sp->d_ws); cudaFree(sp->d_key); cudaFree(sp->d_err);
    delete sp;
    return 0;
}

int pgk_nvf4_pack_rows(const void* data_kn, const void* scale_kn, void* packed, int K, int N, int ld, int col0, int group,
                       int stride, void* stream) {
    PGK_REQUIRE(K >= 32 && K % 32 == 0 && N >= 1 && group >= 1 && ld >= 1, "nvf4_pack_rows failed: invalid shape K=%d N=%d", K, N);
    const size_t total = (size_t)N * (K / 2 + K / 32);
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, 148 * 16);
    PGK_CUDA(launch(nvf4_pack_rows_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const uint8_t*)data_kn,
                    (const uint8_t*)scale_kn, (uint8_t*)packed, K, N, ld, col0, group, stride));
    return 0;
}

}  // extern "C"
