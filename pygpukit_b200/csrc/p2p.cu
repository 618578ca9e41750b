// p2p.cu -- one-shot all-reduce (+ residual add) over NVLink peer memory for the row-parallel layers of the
// tensor-parallel decode step (net-new: the reference is single-GPU; SURVEY.md section 8e names this as the
// replacement of the NCCL all-reduce, whose ~10 us latency per 16 KB message dominates a batch-1 step).
//
//   hidden[M,H] <- bf16( hidden + bf16( sum_r partial_r[M,H] ) )        r = 0 .. world-1, fixed order
//
// One process per GPU.  Every rank owns a communication buffer (cudaMalloc + cudaIpc handle, opened by all
// peers): two data slabs (parity of the call sequence number) and a flag word per (parity, CTA, source rank).
// A call = one kernel, PUSH style (no NVLink round trip on the critical path: remote stores are posted, every
// load is local):
//   1. every CTA stores its slice of the local partial sums into the region of THIS rank in EVERY rank's slab
//      (remote NVLink stores; own buffer included),
//   2. one thread per destination publishes: __threadfence_system, then stores the call's sequence number into
//      the flag word of THIS rank in that peer's buffer,
//   3. `world` threads spin on the LOCAL flag words until every source has published the same number,
//   4. every thread sums its 16-byte vectors over all source regions of the LOCAL slab in rank order (f32
//      accumulate), rounds to bf16, adds the residual and stores locally.
// All ranks compute the identical sum in the identical order, so the replicated hidden state stays bit-identical
// across ranks (greedy tokens cannot diverge between ranks).  Slab reuse is safe with two parities: a rank can
// only reach call k+2 after every rank has published call k+1, i.e. after every rank finished reading call k.
// Sequence numbers live in device memory, so the kernel is CUDA-graph replayable.
#include "common.cuh"
#include "gemv_common.cuh"
#include "../../include/pgk_b200.h"

#include <cstring>

namespace pgk {

constexpr int kP2pMaxWorld = 8;
constexpr int kP2pMaxCtas = 64;
constexpr size_t kP2pSlabBytes = 8u << 20;                  // per parity: one region per source rank
constexpr size_t kP2pRegionBytes = kP2pSlabBytes / kP2pMaxWorld;   // 1 MB: up to 64 rows x 8192 bf16
constexpr size_t kP2pFlagBytes = 2 * kP2pMaxCtas * kP2pMaxWorld * sizeof(unsigned);
// LL ("low latency") region used by the one-launch step kernel (step.cu): per parity and source rank 8192 words of
// 8 bytes {payload, sequence number}; a word is valid when its sequence number matches, so no fence or flag is needed
constexpr size_t kP2pLLWords = 8192;
constexpr size_t kP2pLLBytes = 2 * kP2pMaxWorld * kP2pLLWords * 8;
constexpr size_t kP2pLLOffset = 2 * kP2pSlabBytes + kP2pFlagBytes;
constexpr size_t kP2pBufBytes = kP2pLLOffset + kP2pLLBytes;

struct P2pState {
    int rank = 0, world = 1;
    unsigned char* local = nullptr;                  // this rank's buffer
    unsigned char* peer[kP2pMaxWorld] = {};          // every rank's buffer as mapped here (peer[rank] == local)
    unsigned* seq = nullptr;                         // [kP2pMaxCtas] call counters (device)
    unsigned* err = nullptr;                         // set when a spin timed out
    bool ready = false;
};
static P2pState g_p2p;

struct P2pParams {
    const uint16_t* partial;     // [n] local partial sums (bf16)
    uint16_t* out;               // [n] hidden, updated in place (residual == out) or plain output
    const uint16_t* residual;    // may be null
    unsigned char* peer[kP2pMaxWorld];
    unsigned* seq; unsigned* err;
    int rank, world;
    long n;                      // elements, multiple of 8
    long per_cta;                // elements per CTA, multiple of 8
};

__device__ __forceinline__ uint4 ld_peer16(const void* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_flag_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_flag_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(1024) p2p_allreduce_add_kernel(P2pParams p) {
    __shared__ unsigned s_seq;
    const int cta = blockIdx.x, tid = threadIdx.x;
    // Before the dependency wait: the call counter (internal state, last written by the previous all-reduce,
    // i.e. >= 2 kernels back).  One 16-byte vector per thread: a CTA covers at most 8192 elements.
    if (tid == 0) s_seq = ++p.seq[cta];               // same count on every rank: SPMD, every rank runs every call
    // CTA c always owns elements [c * 8192, (c + 1) * 8192): a slab byte is only ever touched by the same CTA index on
    // every rank, so the per-CTA sequence / parity argument holds whatever the size of the previous call was
    const long e0 = (long)cta * p.per_cta, e1 = min(p.n, e0 + p.per_cta);
    const long e = e0 + (long)tid * 8;
    const bool active = e < e1;
    griddep_wait();                                   // the partial sums come from the GEMV just before
    griddep_launch();
    uint4 rv = make_uint4(0u, 0u, 0u, 0u);            // residual: requested now, consumed after the flag wait
    if (active && p.residual) rv = *reinterpret_cast<const uint4*>(p.residual + e);
    __syncthreads();
    const unsigned seq = s_seq;
    const unsigned par = seq & 1u;
    unsigned char* mine = p.peer[p.rank];
    // 1. local partials -> this rank's region in every rank's slab (posted remote stores)
    const size_t my_region = par * kP2pSlabBytes + (size_t)p.rank * kP2pRegionBytes;
    if (active) {
        const uint4 v = *reinterpret_cast<const uint4*>(p.partial + e);
#pragma unroll
        for (int r = 0; r < kP2pMaxWorld; ++r)
            if (r < p.world) *reinterpret_cast<uint4*>(p.peer[r] + my_region + e * 2) = v;
    }
    __syncthreads();
    // 2. publish to every rank (own flag word included), 3. wait for every source rank's slice
    if (tid < p.world) {
        __threadfence_system();
        unsigned* f = reinterpret_cast<unsigned*>(p.peer[tid] + 2 * kP2pSlabBytes) +
                      ((size_t)par * kP2pMaxCtas + cta) * kP2pMaxWorld + p.rank;
        st_flag_sys(f, seq);
        const unsigned* g = reinterpret_cast<const unsigned*>(mine + 2 * kP2pSlabBytes) +
                            ((size_t)par * kP2pMaxCtas + cta) * kP2pMaxWorld + tid;
        const long long t0 = clock64();
        while (ld_flag_sys(g) != seq) {
            if (clock64() - t0 > 4000000000LL) { *p.err = 1u; break; }     // ~2 s: a peer is gone; do not hang the GPU
        }
    }
    __syncthreads();
    // 4. sum the source regions of the local slab in rank order, round, add the residual
    if (active) {
        uint4 pv[kP2pMaxWorld];
#pragma unroll
        for (int r = 0; r < kP2pMaxWorld; ++r)
            if (r < p.world) pv[r] = ld_peer16(mine + par * kP2pSlabBytes + (size_t)r * kP2pRegionBytes + e * 2);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
        for (int r = 0; r < kP2pMaxWorld; ++r) {
            if (r < p.world) {
                const uint4 v = pv[r];
                acc[0] += bf16lo(v.x); acc[1] += bf16hi(v.x); acc[2] += bf16lo(v.y); acc[3] += bf16hi(v.y);
                acc[4] += bf16lo(v.z); acc[5] += bf16hi(v.z); acc[6] += bf16lo(v.w); acc[7] += bf16hi(v.w);
            }
        }
        uint16_t o[8];
        if (p.residual) {
            const uint16_t* rb = reinterpret_cast<const uint16_t*>(&rv);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = hadd_bf16_bits(acc[j], rb[j]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = f32_to_bf16_bits(acc[j]);
        }
        *reinterpret_cast<uint4*>(p.out + e) = *reinterpret_cast<const uint4*>(o);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// LL ("low latency") form for one decode row (n <= 8192): every 8-byte word carries {two bf16 partial sums, exchange
// number}.  An aligned 8-byte store is delivered atomically, so a word whose number matches is valid: no system fence,
// no separate flag, the critical path is ONE NVLink hop (the flag form above pays a fence round trip before its flag).
// Exchange numbers come from one device word per rank (shared with the step kernel, step.cu) that every rank advances
// identically (SPMD).  One CTA; thread t owns elements 8t .. 8t+7 = words 4t .. 4t+3 of its rank's region.
struct P2pLLParams {
    const uint16_t* partial; uint16_t* out; const uint16_t* residual;
    unsigned char* ll[kP2pMaxWorld];
    unsigned* epoch; unsigned* err;
    int rank, world, n;
};
__device__ __forceinline__ void st_ll16(unsigned char* p, uint32_t d0, uint32_t d1, uint32_t seq) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(d0), "r"(seq), "r"(d1), "r"(seq) : "memory");
}
__device__ __forceinline__ uint4 ld_ll16(const unsigned char* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__global__ void __launch_bounds__(1024) p2p_ll_allreduce_add_kernel(P2pLLParams p) {
    const int tid = threadIdx.x;
    const int e = tid * 8;
    const bool active = e < p.n;
    griddep_wait();                                   // the partial sums come from the GEMV just before
    griddep_launch();
    const unsigned seq = *p.epoch + 1u;
    const size_t par_off = (size_t)(seq & 1u) * kP2pMaxWorld * kP2pLLWords * 8;
    if (active) {
        const uint4 v = *reinterpret_cast<const uint4*>(p.partial + e);
        uint4 rv = make_uint4(0u, 0u, 0u, 0u);
        if (p.residual) rv = *reinterpret_cast<const uint4*>(p.residual + e);
        const size_t mine = par_off + ((size_t)p.rank * kP2pLLWords + (size_t)tid * 4) * 8;
#pragma unroll
        for (int r = 0; r < kP2pMaxWorld; ++r) {
            if (r < p.world) {
                st_ll16(p.ll[r] + mine, v.x, v.y, seq);
                st_ll16(p.ll[r] + mine + 16, v.z, v.w, seq);
            }
        }
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        const unsigned char* base = p.ll[p.rank] + par_off + (size_t)tid * 32;
        const long long t0 = clock64();
#pragma unroll 1
        for (int r = 0; r < p.world; ++r) {
            const unsigned char* src = base + (size_t)r * kP2pLLWords * 8;
            uint4 a = ld_ll16(src), b = ld_ll16(src + 16);
            while (a.y != seq || a.w != seq || b.y != seq || b.w != seq) {
                if (clock64() - t0 > 4000000000LL) { *p.err = 1u; break; }       // ~2 s: a peer is gone; do not hang the GPU
                a = ld_ll16(src); b = ld_ll16(src + 16);
            }
            acc[0] += bf16lo(a.x); acc[1] += bf16hi(a.x); acc[2] += bf16lo(a.z); acc[3] += bf16hi(a.z);
            acc[4] += bf16lo(b.x); acc[5] += bf16hi(b.x); acc[6] += bf16lo(b.z); acc[7] += bf16hi(b.z);
        }
        uint16_t o[8];
        if (p.residual) {
            const uint16_t* rb = reinterpret_cast<const uint16_t*>(&rv);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = hadd_bf16_bits(acc[j], rb[j]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = f32_to_bf16_bits(acc[j]);
        }
        *reinterpret_cast<uint4*>(p.out + e) = *reinterpret_cast<const uint4*>(o);
    }
    __syncthreads();
    if (tid == 0) *p.epoch = seq;
}

// greedy token of a vocabulary-sharded lm_head: every rank publishes (its best logit, global index) as two LL words and
// takes the maximum of all keys (identical on every rank; the lowest index wins ties) -- 16 bytes per rank instead of
// an all-gather of the logits
__global__ void p2p_ll_argmax_kernel(const uint16_t* __restrict__ logits, int32_t* token_buf, int vocab_offset, P2pLLParams p) {
    griddep_wait();
    griddep_launch();
    if (threadIdx.x != 0) return;
    const unsigned seq = *p.epoch + 1u;
    const size_t par_off = (size_t)(seq & 1u) * kP2pMaxWorld * kP2pLLWords * 8;
    const int idx = token_buf[0];
    const unsigned long long k = argmax_key(bf16_bits_to_f32(logits[idx]), idx + vocab_offset);
    for (int r = 0; r < p.world; ++r)
        st_ll16(p.ll[r] + par_off + (size_t)p.rank * kP2pLLWords * 8, (uint32_t)(k & 0xffffffffull), (uint32_t)(k >> 32), seq);
    unsigned long long best = 0ull;
    const long long t0 = clock64();
    for (int r = 0; r < p.world; ++r) {
        const unsigned char* src = p.ll[p.rank] + par_off + (size_t)r * kP2pLLWords * 8;
        uint4 a = ld_ll16(src);
        while (a.y != seq || a.w != seq) {
            if (clock64() - t0 > 4000000000LL) { *p.err = 1u; break; }
            a = ld_ll16(src);
        }
        const unsigned long long kr = ((unsigned long long)a.z << 32) | a.x;
        best = kr > best ? kr : best;
    }
    token_buf[0] = 0x7fffffff - (int)(uint32_t)(best & 0xffffffffull);
    *p.epoch = seq;
}

// step.cu: the LL regions of every rank as mapped here, and the device word that numbers the exchanges (every rank runs
// the same programs in the same order, so the numbering agrees without communication)
bool p2p_ll_info(int* rank, int* world, unsigned char** ll_peers, unsigned** epoch, size_t* words_per_source) {
    if (!g_p2p.ready) return false;
    *rank = g_p2p.rank; *world = g_p2p.world;
    for (int r = 0; r < kP2pMaxWorld; ++r) ll_peers[r] = r < g_p2p.world ? g_p2p.peer[r] + kP2pLLOffset : nullptr;
    *epoch = g_p2p.seq + kP2pMaxCtas + 1;
    *words_per_source = kP2pLLWords;
    return true;
}

// gemv_fp8.cu / gemv_nvf4.cu: the same regions and exchange numbering for the GEMV kernels that do their own all-reduce
bool p2p_gemv_xchg(GemvXchg* x) {
    if (!g_p2p.ready) return false;
    for (int r = 0; r < kP2pMaxWorld; ++r) x->ll[r] = r < g_p2p.world ? g_p2p.peer[r] + kP2pLLOffset : nullptr;
    x->epoch = g_p2p.seq + kP2pMaxCtas + 1;
    x->done = g_p2p.seq + kP2pMaxCtas + 2;
    x->err = g_p2p.err;
    x->rank = g_p2p.rank; x->world = g_p2p.world;
    x->slot_words = (unsigned)kP2pLLWords;
    x->par_bytes = (unsigned)(kP2pMaxWorld * kP2pLLWords * 8);
    return true;
}

}  // namespace pgk

using namespace pgk;

extern "C" {

size_t pgk_p2p_handle_bytes(void) { return sizeof(cudaIpcMemHandle_t); }

/* step 1 (every rank): allocate the local buffer, return its IPC handle (host memory, pgk_p2p_handle_bytes()) */
int pgk_p2p_create(int rank, int world, void* host_handle_out) {
    PGK_REQUIRE(world >= 1 && world <= kP2pMaxWorld && rank >= 0 && rank < world, "p2p_create failed: bad rank %d / world %d", rank, world);
    PGK_REQUIRE(!g_p2p.local, "p2p_create failed: already created");
    g_p2p.rank = rank; g_p2p.world = world;
    PGK_CUDA(cudaMalloc(&g_p2p.local, kP2pBufBytes));
    PGK_CUDA(cudaMemset(g_p2p.local, 0, kP2pBufBytes));
    PGK_CUDA(cudaMalloc(&g_p2p.seq, (kP2pMaxCtas + 3) * sizeof(unsigned)));
    PGK_CUDA(cudaMemset(g_p2p.seq, 0, (kP2pMaxCtas + 3) * sizeof(unsigned)));
    g_p2p.err = g_p2p.seq + kP2pMaxCtas;
    cudaIpcMemHandle_t h;
    PGK_CUDA(cudaIpcGetMemHandle(&h, g_p2p.local));
    memcpy(host_handle_out, &h, sizeof(h));
    PGK_CUDA(cudaDeviceSynchronize());
    return 0;
}

/* step 2 (every rank, after an all-gather of the handles over the host channel): map every peer's buffer */
int pgk_p2p_open(const void* host_handles_all) {
    PGK_REQUIRE(g_p2p.local && !g_p2p.ready, "p2p_open failed: call p2p_create first (once)");
    const cudaIpcMemHandle_t* hs = (const cudaIpcMemHandle_t*)host_handles_all;
    for (int r = 0; r < g_p2p.world; ++r) {
        if (r == g_p2p.rank) { g_p2p.peer[r] = g_p2p.local; continue; }
        void* ptr = nullptr;
        PGK_CUDA(cudaIpcOpenMemHandle(&ptr, hs[r], cudaIpcMemLazyEnablePeerAccess));
        g_p2p.peer[r] = (unsigned char*)ptr;
    }
    g_p2p.ready = true;
    return 0;
}

int pgk_p2p_ready(void) { return g_p2p.ready ? 1 : 0; }
/* largest message (bytes) pgk_p2p_allreduce_add_bf16 takes; larger ones go through NCCL */
size_t pgk_p2p_max_bytes(void) { return kP2pRegionBytes; }

/* out[n] = bf16(residual + bf16(sum over ranks of partial)) (residual NULL: plain all-reduce); out may alias
 * residual.  n % 8 == 0, n * 2 bytes <= 1 MB, 16-byte aligned pointers.  Every rank must make the same calls
 * in the same order. */
int pgk_p2p_allreduce_add_bf16(const void* partial, const void* residual, void* out, size_t n, void* stream) {
    PGK_REQUIRE(g_p2p.ready, "p2p_allreduce failed: peer buffers not opened (pgk_p2p_create / pgk_p2p_open)");
    PGK_REQUIRE(n % 8 == 0 && n * 2 <= kP2pRegionBytes, "p2p_allreduce failed: n = %zu must be a multiple of 8 and <= %zu", n, kP2pRegionBytes / 2);
    PGK_REQUIRE(((uintptr_t)partial % 16 == 0) && ((uintptr_t)out % 16 == 0) && ((uintptr_t)residual % 16 == 0),
                "p2p_allreduce failed: pointers must be 16-byte aligned");
    P2pParams p = {};
    p.partial = (const uint16_t*)partial; p.out = (uint16_t*)out; p.residual = (const uint16_t*)residual;
    for (int r = 0; r < g_p2p.world; ++r) p.peer[r] = g_p2p.peer[r];
    p.seq = g_p2p.seq; p.err = g_p2p.err; p.rank = g_p2p.rank; p.world = g_p2p.world;
    p.n = (long)n;
    // one CTA of 1024 threads per 8192 elements (a fixed 16 KB slot per CTA index): a batch-1 row of H <= 8192 is a
    // single CTA, 1 MB is 64 CTAs
    const int ctas = (int)((n + 8191) / 8192);
    p.per_cta = 8192;
    PGK_REQUIRE(ctas <= kP2pMaxCtas, "p2p_allreduce failed: %d CTAs (internal limit %d)", ctas, kP2pMaxCtas);
    PGK_CUDA(launch(p2p_allreduce_add_kernel, dim3(ctas), dim3(1024), 0, (cudaStream_t)stream, p));
    return 0;
}

/* one decode row (n <= 8192, n % 8 == 0): the LL form of pgk_p2p_allreduce_add_bf16 -- same result, one NVLink hop */
int pgk_p2p_ll_allreduce_add_bf16(const void* partial, const void* residual, void* out, size_t n, void* stream) {
    PGK_REQUIRE(g_p2p.ready, "p2p_ll_allreduce failed: peer buffers not opened (pgk_p2p_create / pgk_p2p_open)");
    PGK_REQUIRE(n % 8 == 0 && n >= 8 && n <= 8192, "p2p_ll_allreduce failed: n = %zu must be a multiple of 8 and <= 8192", n);
    PGK_REQUIRE(((uintptr_t)partial % 16 == 0) && ((uintptr_t)out % 16 == 0) && ((uintptr_t)residual % 16 == 0),
                "p2p_ll_allreduce failed: pointers must be 16-byte aligned");
    P2pLLParams p = {};
    p.partial = (const uint16_t*)partial; p.out = (uint16_t*)out; p.residual = (const uint16_t*)residual;
    for (int r = 0; r < g_p2p.world; ++r) p.ll[r] = g_p2p.peer[r] + kP2pLLOffset;
    p.epoch = g_p2p.seq + kP2pMaxCtas + 1; p.err = g_p2p.err; p.rank = g_p2p.rank; p.world = g_p2p.world; p.n = (int)n;
    const int threads = (int)(((n / 8) + 31) / 32 * 32);
    PGK_CUDA(launch(p2p_ll_allreduce_add_kernel, dim3(1), dim3(threads), 0, (cudaStream_t)stream, p));
    return 0;
}

/* token_buf[0] holds this rank's LOCAL argmax index over logits_local (bf16 [V_local]); on return it holds the global
 * greedy token (identical on every rank).  Replaces the all-gather of the logits + argmax over world * V_local. */
int pgk_p2p_ll_argmax_bf16(const void* logits_local, int32_t* token_buf, int vocab_offset, void* stream) {
    PGK_REQUIRE(g_p2p.ready, "p2p_ll_argmax failed: peer buffers not opened (pgk_p2p_create / pgk_p2p_open)");
    P2pLLParams p = {};
    for (int r = 0; r < g_p2p.world; ++r) p.ll[r] = g_p2p.peer[r] + kP2pLLOffset;
    p.epoch = g_p2p.seq + kP2pMaxCtas + 1; p.err = g_p2p.err; p.rank = g_p2p.rank; p.world = g_p2p.world;
    PGK_CUDA(launch(p2p_ll_argmax_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, (const uint16_t*)logits_local, token_buf,
                    vocab_offset, p));
    return 0;
}

/* 1 if a wait timed out since creation (a peer died): results after that are undefined */
int pgk_p2p_error(int* host_flag) {
    unsigned v = 0;
    if (g_p2p.err) PGK_CUDA(cudaMemcpy(&v, g_p2p.err, sizeof(v), cudaMemcpyDeviceToHost));
    *host_flag = (int)v;
    return 0;
}

int pgk_p2p_destroy(void) {
    if (!g_p2p.local) return 0;
    cudaDeviceSynchronize();
    for (int r = 0; r < g_p2p.world; ++r)
        if (r != g_p2p.rank && g_p2p.peer[r]) cudaIpcCloseMemHandle(g_p2p.peer[r]);
    cudaFree(g_p2p.local);
    cudaFree(g_p2p.seq);
    g_p2p = P2pState();
    return 0;
}

}  // extern "C"
