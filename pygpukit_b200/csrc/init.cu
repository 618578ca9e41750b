// init.cu -- device-side synthetic weight initialisation and FP8 block quantisation.
//
// The benchmarks run random-init models of 8B..70B parameters; generating and quantising those on
// the host (NumPy) takes minutes per model, so the N(0, std^2) BF16 masters are drawn on the device
// (counter-based hash RNG + Box-Muller) and quantised there:
//   * FP8: 128x128 block absmax / 448 -> bf16 scale_inv, codes = round-to-nearest-even E4M3 of
//     w / scale (saturating at +-448, never 0x7F/0xFF).  Same definition as the host quantiser
//     pygpukit_b200/llm/quant.py:quantize_fp8_e4m3_block (the reference has no FP8 quantiser; it only
//     reads checkpoints: llm/loader.py:197-252, llm/quant.py:30-57).
//   * NVF4: pgk_quantize_bf16_to_nvf4 (quant_nvf4.cu, bit-exact with the reference).
// Also the one-thread "advance" kernel that bumps the device-side position / context counters
// between graph replays.
#include "common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

__global__ void fill_normal_bf16_kernel(uint16_t* __restrict__ out, size_t n, uint32_t seed, float std) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; 2 * i < n; i += stride) {
        uint32_t lo = (uint32_t)i, hi = (uint32_t)(i >> 32);
        uint32_t a = mix32(lo * 2u + 0x9e3779b9u * (seed + 1u) + hi * 0x85ebca6bu);
        uint32_t b = mix32(a ^ (lo * 2u + 1u) ^ (seed * 0xc2b2ae35u));
        float u1 = ((float)(a >> 8) + 1.0f) * (1.0f / 16777216.0f);   // (0, 1]
        float u2 = (float)(b >> 8) * (1.0f / 16777216.0f);            // [0, 1)
        float r = sqrtf(-2.0f * __logf(u1)) * std;
        float s, c;
        __sincosf(6.283185307179586f * u2, &s, &c);
        out[2 * i] = f32_to_bf16_bits(r * c);
        if (2 * i + 1 < n) out[2 * i + 1] = f32_to_bf16_bits(r * s);
    }
}

// one CTA (256 threads) per 128x128 block of W[N,K] (bf16)
__global__ void __launch_bounds__(256)
quantize_fp8_block_kernel(const uint16_t* __restrict__ W, uint8_t* __restrict__ codes,
                          uint16_t* __restrict__ scale, int N, int K) {
    __shared__ float red[8];
    __shared__ float s_scale;
    const int bk = blockIdx.x, bn = blockIdx.y;
    const int SK = (K + 127) >> 7;
    const int tid = threadIdx.x;
    float mx = 0.f;
    for (int i = tid; i < 128 * 128; i += 256) {
        int n = bn * 128 + (i >> 7), k = bk * 128 + (i & 127);
        if (n < N && k < K) mx = fmaxf(mx, fabsf(bf16_bits_to_f32(W[(size_t)n * K + k])));
    }
    mx = warp_max(mx);
    if ((tid & 31) == 0) red[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        float m = red[0];
        for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
        float sc = m > 0.f ? __fdiv_rn(m, 448.0f) : 1.0f;
        uint16_t sb = f32_to_bf16_bits(sc);
        scale[(size_t)bn * SK + bk] = sb;
        s_scale = bf16_bits_to_f32(sb);
    }
    __syncthreads();
    const float sc = s_scale;
    for (int i = tid; i < 128 * 128; i += 256) {
        int n = bn * 128 + (i >> 7), k = bk * 128 + (i & 127);
        if (n < N && k < K) {
            float v = __fdiv_rn(bf16_bits_to_f32(W[(size_t)n * K + k]), sc);
            uint16_t pk;
            asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(pk) : "f"(0.0f), "f"(v));
            codes[(size_t)n * K + k] = (uint8_t)(pk & 0xff);
        }
    }
}

__global__ void decode_advance_kernel(int32_t* a, int32_t* b, int delta, int n) {
    griddep_wait();
    griddep_launch();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (a) a[i] += delta;
        if (b) b[i] += delta;
    }
}

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_fill_normal_bf16(void* out, size_t n, uint32_t seed, float std, void* stream) {
    if (n == 0) return 0;
    size_t pairs = (n + 1) / 2;
    unsigned grid = (unsigned)((pairs + 255) / 256);
    unsigned cap = (unsigned)sm_count() * 16u;
    if (grid > cap) grid = cap;
    PGK_CUDA(launch(fill_normal_bf16_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (uint16_t*)out, n, seed, std));
    return 0;
}

int pgk_quantize_bf16_to_fp8_block(const void* w_nk, void* codes, void* scale, int N, int K, void* stream) {
    PGK_REQUIRE(N >= 1 && K >= 1, "quantize_bf16_to_fp8_block failed: invalid shape N=%d K=%d", N, K);
    dim3 grid((K + 127) / 128, (N + 127) / 128);
    PGK_CUDA(launch(quantize_fp8_block_kernel, grid, dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)w_nk,
                    (uint8_t*)codes, (uint16_t*)scale, N, K));
    return 0;
}

int pgk_decode_advance(int32_t* pos_buf, int32_t* ctx_buf, int delta, void* stream) {
    PGK_CUDA(launch(decode_advance_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, pos_buf, ctx_buf, delta, 1));
    return 0;
}
/* the same for n sequences (arrays of n positions / contexts) */
int pgk_decode_advance_n(int32_t* pos_buf, int32_t* ctx_buf, int delta, int n, void* stream) {
    PGK_CUDA(launch(decode_advance_kernel, dim3(1), dim3(64), 0, (cudaStream_t)stream, pos_buf, ctx_buf, delta, n));
    return 0;
}

}  // extern "C"
