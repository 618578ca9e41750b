// common.cuh -- shared device/host helpers for libpgk_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <string>

namespace pgk {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
int  cuda_fail(cudaError_t e, const char* what, const char* file, int line);
void count_launch(int n = 1);
bool pdl_enabled();
int  sm_count();

// E4M3 codes 0x7F / 0xFF: +-480 in the reference's CUDA tables (fp8.cuh:89,121) but NaN for the hardware conversion
// (cvt.rn.f16x2.e4m3x2) the streaming kernels use.  Every FP8 weight tensor is scanned ONCE, at its first use
// (outside graph capture), and remembered by base pointer: tensors that hold such codes are served by the exact
// table-semantics kernel, so one entry point gives one answer.  pgk_free() forgets the pointer.
int  fp8_codes_clean(const void* codes, size_t nbytes, cudaStream_t st, bool* clean);
// A row-parallel GEMV and the all-reduce after it in ONE launch (tensor parallel decode): the thread that finishes output
// element n publishes its bf16 partial to every rank as an 8-byte LL word {bits, exchange number} over peer memory and
// sums the words of all ranks in rank order (+ residual) -- the arithmetic of p2p_ll_allreduce_add_kernel (p2p.cu),
// without its launch.  world == 0: no exchange.
struct GemvXchg {
    unsigned char* ll[8];     // LL region of every rank (peer-mapped pointers)
    unsigned* epoch;          // exchange number of the last completed exchange (device word, advanced identically on every rank)
    unsigned* done;           // arrival counter of the CTAs that take part in the exchange (zero between launches)
    unsigned* err;            // set when a wait timed out (a peer is gone)
    int rank, world;
    unsigned slot_words;      // words per source rank in a region
    unsigned par_bytes;       // bytes between the two parity halves of a region
};
bool p2p_gemv_xchg(GemvXchg* x);     // p2p.cu: false until the peer buffers are open
void fp8_codes_forget(const void* base);


#define PGK_CUDA(call)                                                        \
    do {                                                                      \
        cudaError_t _e = (call);                                              \
        if (_e != cudaSuccess) return pgk::cuda_fail(_e, #call, __FILE__, __LINE__); \
    } while (0)

#define PGK_REQUIRE(cond, ...)                                                \
    do {                                                                      \
        if (!(cond)) { pgk::set_error(__VA_ARGS__); return 1; }               \
    } while (0)

#define PGK_LAUNCH_CHECK(name)                                                \
    do {                                                                      \
        cudaError_t _e = cudaGetLastError();                                  \
        if (_e != cudaSuccess) return pgk::cuda_fail(_e, name, __FILE__, __LINE__); \
    } while (0)

// Launch with optional programmatic-dependent-launch attribute (PDL): the kernel may
// start while its predecessor drains; every kernel calls griddep_wait() before it
// touches data produced by an earlier kernel.
// One shared-memory / L1 split for every kernel of the step (PGK_CARVEOUT, percent of the maximum; -1 = driver's choice)
inline int smem_carveout() {
    static const int v = [] { const char* e = getenv("PGK_CARVEOUT"); return e ? atoi(e) : -1; }();
    return v;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                          cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int nattr = 0;
    if (pdl_enabled()) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        nattr = 1;
    }
    if (const int co = smem_carveout(); co >= 0) {
        attr[nattr].id = cudaLaunchAttributePreferredSharedMemoryCarveout;
        attr[nattr].val.sharedMemCarveout = (unsigned)co;
        ++nattr;
    }
    cfg.attrs = attr;
    cfg.numAttrs = nattr;
    count_launch();
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__

// PDL discipline of this library: every kernel executes griddepcontrol.wait BEFORE launch_dependents.  A
// dependent grid can therefore only start once its predecessor has passed its own wait, i.e. once every
// kernel before the predecessor has completed and flushed: code placed before griddep_wait() may read
// anything that was last written two or more kernels back (weights, the device-side position counters,
// KV-cache rows of earlier steps), never the predecessor's outputs.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float bf16_bits_to_f32(uint32_t h) { return __uint_as_float(h << 16); }
__device__ __forceinline__ float bf16lo(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float bf16hi(uint32_t p) { return __uint_as_float(p & 0xffff0000u); }
__device__ __forceinline__ uint16_t f32_to_bf16_bits(float f) {
    return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}
__device__ __forceinline__ float round_bf16(float f) {
    return __bfloat162float(__float2bfloat16_rn(f));
}
// bf16(a) + b with ONE rounding (native bf16 add = the reference's __hadd in
// add_inplace_bf16_kernel, elementwise_kernels.cuh:381-390); returns bf16 bits.
__device__ __forceinline__ uint16_t hadd_bf16_bits(float a, uint16_t b_bits) {
    __nv_bfloat16 r = __hadd(__float2bfloat16_rn(a), __ushort_as_bfloat16(b_bits));
    return __bfloat16_as_ushort(r);
}
// two f32 -> packed bf16x2 with ONE full-rate instruction (cvt.rn.bf16x2.f32 = F2FP.BF16.F32.PACK_AB); the
// scalar F2F.BF16.F32 runs on the quarter-rate conversion pipe
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// streaming loads: read-only path, do not allocate in L1, evict-first in L2 (weights are
// touched once per token).  sm_100a has 256-bit global loads (LDG.E.NA.EFL2.256).
struct alignas(32) u32x8 { uint32_t v[8]; };
__device__ __forceinline__ u32x8 ldg_stream256(const void* p) {
    u32x8 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]),
                   "=r"(r.v[6]), "=r"(r.v[7]) : "l"(p));
    return r;
}
// activations written by the kernel just before this one (programmatic dependent launch: this kernel's lifetime
// overlaps its producer's, so the read-only .nc path is formally off limits for them): coherent, L2-cached loads
__device__ __forceinline__ uint4 ld_act16(const void* p) {
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint16_t ld_act2(const uint16_t* p) {
    uint16_t v;
    asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// acc += a(f16) * b(f16), product exact, accumulate in f32.  One FHFMA on sm_100a.
// lo / hi select the 16-bit half of the packed registers.
__device__ __forceinline__ void fhfma_f16_lo(float& acc, uint32_t a, uint32_t b) {
    asm("{.reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;"
        " fma.rn.f32.f16 %0, al, bl, %0;}" : "+f"(acc) : "r"(a), "r"(b));
}
__device__ __forceinline__ void fhfma_f16_hi(float& acc, uint32_t a, uint32_t b) {
    asm("{.reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;"
        " fma.rn.f32.f16 %0, ah, bh, %0;}" : "+f"(acc) : "r"(a), "r"(b));
}
__device__ __forceinline__ void fhfma_bf16_lo(float& acc, uint32_t a, uint32_t b) {
    asm("{.reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;"
        " fma.rn.f32.bf16 %0, al, bl, %0;}" : "+f"(acc) : "r"(a), "r"(b));
}
__device__ __forceinline__ void fhfma_bf16_hi(float& acc, uint32_t a, uint32_t b) {
    asm("{.reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;"
        " fma.rn.f32.bf16 %0, ah, bh, %0;}" : "+f"(acc) : "r"(a), "r"(b));
}
// a-half selectable independently: acc += a.lo * b.hi etc.
__device__ __forceinline__ void fhfma_f16_lh(float& acc, uint32_t a, uint32_t b) {
    asm("{.reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;"
        " fma.rn.f32.f16 %0, al, bh, %0;}" : "+f"(acc) : "r"(a), "r"(b));
}
__device__ __forceinline__ void fhfma_f16_hl(float& acc, uint32_t a, uint32_t b) {
    asm("{.reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;"
        " fma.rn.f32.f16 %0, ah, bl, %0;}" : "+f"(acc) : "r"(a), "r"(b));
}

// two e4m3 bytes (low / high half of a 32-bit word) -> f16x2 (exact).  F2FP.F16.E4M3.UNPACK_B
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2_lo(uint32_t w) {
    uint32_t r;
    asm("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.rn.f16x2.e4m3x2 %0, lo;}" : "=r"(r) : "r"(w));
    return r;
}
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2_hi(uint32_t w) {
    uint32_t r;
    asm("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.rn.f16x2.e4m3x2 %0, hi;}" : "=r"(r) : "r"(w));
    return r;
}
// one byte holding two e2m1 nibbles (low nibble -> low half) -> f16x2 (exact).
template <int BYTE>
__device__ __forceinline__ uint32_t e2m1x2_to_f16x2(uint32_t w) {
    uint32_t r;
    asm("{.reg .b8 b0, b1, b2, b3; mov.b32 {b0, b1, b2, b3}, %1; cvt.rn.f16x2.e2m1x2 %0, b%2;}"
        : "=r"(r) : "r"(w), "n"(BYTE));
    return r;
}

// FP8 E4M3 decode exactly as the reference CUDA LUT (fp8.cuh:57-122): 0x7F/0xFF -> +-480.
__device__ __forceinline__ float fp8_e4m3_to_f32(uint32_t v) {
    uint32_t e = (v >> 3) & 0xf, m = v & 7;
    float r = (e == 0) ? (float)m * 0.001953125f : __uint_as_float(((e + 120u) << 23) | (m << 20));
    return (v & 0x80) ? -r : r;
}
// UE4M3 scale decode (nvf4.cuh:65-115): exp==0 is NORMAL ((1+m/8)*2^-7), bit 7 ignored.
__device__ __forceinline__ float ue4m3_to_f32(uint32_t v) {
    return __uint_as_float(((v & 0x7fu) << 20) + 0x3C000000u);
}
__device__ __forceinline__ float e2m1_to_f32(uint32_t nib) {
    const float t[8] = {0.f, 0.5f, 1.f, 1.5f, 2.f, 3.f, 4.f, 6.f};
    float r = t[nib & 7];
    return (nib & 8) ? -r : r;
}


// ---- mbarrier + TMA bulk copy (cp.async.bulk, SASS UBLKCP): contiguous global -> shared, completion
// counted in bytes on an mbarrier.  No tensor map needed for 1-D tiles.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
#endif  // __CUDACC__

}  // namespace pgk
