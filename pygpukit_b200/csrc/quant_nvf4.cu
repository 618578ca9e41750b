// quant_nvf4.cu -- BF16 [K,N] -> PyGPUkit "NVF4" (E2M1 nibbles + UE4M3 scale per 32-k block).
// Defines the weight bytes every NVF4 kernel and the oracle see, so it must be bit-exact
// with quantize_bf16_to_nvf4_kernel (native/ops/matmul/gemv/w4a16_bf16/sm120/
// nvf4_kernels.cu:239-324).  THIS FILE IS COMPILED WITH --use_fast_math, like the reference
// build (native/CMakeLists.txt:122), because `max_abs / 6.0f` and `1.0f / scale` are then
// approximate divisions and the nibble thresholds are compared against their results.
//
// Work split: one thread per (column n, scale block sb) -- coalesced over n for the
// N-contiguous input and outputs.
#include "common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

__device__ __forceinline__ uint32_t snap_e2m1(float v) {
    // nearest-of-8 magnitudes {0,.5,1,1.5,2,3,4,6} with strict '<' thresholds (nvf4_kernels.cu:303-314)
    uint32_t s = (v < 0.f) ? 8u : 0u;
    float a = fabsf(v);
    uint32_t q = (a < 0.25f) ? 0u : (a < 0.75f) ? 1u : (a < 1.25f) ? 2u : (a < 1.75f) ? 3u
               : (a < 2.5f) ? 4u : (a < 3.5f) ? 5u : (a < 5.0f) ? 6u : 7u;
    return s | q;
}

__global__ void quantize_nvf4_kernel(const uint16_t* __restrict__ in, uint8_t* __restrict__ data,
                                     uint8_t* __restrict__ scale_out, int K, int N) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const int sb = blockIdx.y;
    if (n >= N) return;
    const int k0 = sb * 32, k1 = min(K, k0 + 32);
    float max_abs = 0.f;
    for (int k = k0; k < k1; ++k) max_abs = fmaxf(max_abs, fabsf(bf16_bits_to_f32(in[(size_t)k * N + n])));
    // UE4M3 encode: loop-normalise into [1,2), exponent in [-7, 8], mantissa rounded with no
    // carry into the exponent (nvf4_kernels.cu:255-296)
    float scale = (max_abs > 1e-8f) ? (max_abs / 6.0f) : 1.0f;
    int e = 0;
    float norm = scale;
    if (norm >= 2.0f) { while (norm >= 2.0f && e < 8) { norm *= 0.5f; ++e; } }
    else if (norm < 1.0f && norm > 1e-8f) { while (norm < 1.0f && e > -7) { norm *= 2.0f; --e; } }
    int mant = __float2int_rn((norm - 1.0f) * 8.0f);
    mant = max(0, min(7, mant));
    int eb = max(0, min(15, e + 7));
    scale_out[(size_t)sb * N + n] = (uint8_t)((eb << 3) | mant);
    // values are divided by the RE-DECODED scale
    float dec = (1.0f + (float)mant / 8.0f) * ldexpf(1.0f, eb - 7);
    float inv = 1.0f / dec;
    for (int k = k0; k < k1; k += 2) {
        float v0 = bf16_bits_to_f32(in[(size_t)k * N + n]) * inv;
        float v1 = (k + 1 < k1) ? bf16_bits_to_f32(in[(size_t)(k + 1) * N + n]) * inv : 0.f;
        data[(size_t)(k >> 1) * N + n] = (uint8_t)((snap_e2m1(v1) << 4) | snap_e2m1(v0));
    }
}

}  // namespace pgk

extern "C" int pgk_quantize_bf16_to_nvf4(const void* in_kn, void* out_data, void* out_scale, int K, int N,
                                         void* stream) {
    PGK_REQUIRE(K > 0 && N > 0 && K % 2 == 0, "quantize_bf16_to_nvf4 failed: K must be even and positive (K=%d N=%d)", K, N);
    dim3 grid((N + 255) / 256, (K + 31) / 32);
    PGK_CUDA(pgk::launch(pgk::quantize_nvf4_kernel, grid, dim3(256), 0, (cudaStream_t)stream,
                         (const uint16_t*)in_kn, (uint8_t*)out_data, (uint8_t*)out_scale, K, N));
    return 0;
}
