// gemm_w8a16.cu -- W8A16 GEMM with the reference's [K,N] weight layout:
//   C[M,N] = A[M,K] . dequant(B[K,N]),  S = bf16 [ceil(K/128), ceil(N/128)]
// replaces pygpukit_w8a16_gemm_sm120 (native/ops/matmul/gemm/w8a16_bf16/sm120/
// w8a16_gemm.cu:104-456): its scalar kernel (M<16) reads B uncoalesced per thread and its
// mma.sync kernel is unpipelined.  Here B is N-contiguous, so lanes own 4 consecutive columns
// (one 32-bit load = 4 codes), K is split over the warps of a CTA and (for skinny M) over
// CTAs, and rows are processed in register tiles of MT activations so the weights are read
// once per tile.  Weight value = LUT[code]*scale as in the reference GEMV/scalar path
// (0x7F/0xFF decode to NaN here, the hardware E4M3 meaning; checkpoints never contain them).
// The llm layers do not use this layout: LinearFP8 keeps [N,K] and calls gemv_fp8 /
// the tcgen05 skinny GEMM; this entry point exists for API parity (ops.w8a16_gemm_sm120).
#include "common.cuh"
#include "../../include/pgk_b200.h"

namespace pgk {

constexpr int kGW = 8;   // warps per CTA (K split)
template <int MT>
__global__ void __launch_bounds__(kGW * 32)
w8a16_gemm_kn_kernel(const uint16_t* __restrict__ A, const uint8_t* __restrict__ B, const uint16_t* __restrict__ S,
                     uint16_t* __restrict__ C, int M, int K, int N) {
    __shared__ float red[kGW][MT][128];
    griddep_wait();
    griddep_launch();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n0 = blockIdx.x * 128 + lane * 4;
    const int m0 = blockIdx.y * MT;
    const int SN = (N + 127) >> 7;
    float acc[MT][4];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[m][j] = 0.f;
    const bool full = (n0 + 3 < N) && ((N & 3) == 0);
    for (int k = warp; k < K; k += kGW) {
        uint32_t w = 0;
        if (full) w = __ldg(reinterpret_cast<const uint32_t*>(B + (size_t)k * N + n0));
        else {
#pragma unroll
            for (int j = 0; j < 4; ++j) if (n0 + j < N) w |= (uint32_t)B[(size_t)k * N + n0 + j] << (8 * j);
        }
        float sc = (n0 < N) ? bf16_bits_to_f32(S[(size_t)(k >> 7) * SN + (n0 >> 7)]) : 0.f;
        float wf[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) wf[j] = fp8_e4m3_to_f32((w >> (8 * j)) & 0xff) * sc;
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            float a = (m0 + m < M) ? bf16_bits_to_f32(A[(size_t)(m0 + m) * K + k]) : 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[m][j] = fmaf(a, wf[j], acc[m][j]);
        }
    }
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int j = 0; j < 4; ++j) red[warp][m][lane * 4 + j] = acc[m][j];
    __syncthreads();
    for (int t = threadIdx.x; t < MT * 128; t += blockDim.x) {
        int m = t / 128, c = t % 128;
        if (m0 + m >= M || blockIdx.x * 128 + c >= N) continue;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kGW; ++w) v += red[w][m][c];
        C[(size_t)(m0 + m) * N + blockIdx.x * 128 + c] = f32_to_bf16_bits(v);
    }
}

}  // namespace pgk

extern "C" int pgk_w8a16_gemm(const void* A, const void* B_kn, const void* S, void* C, int M, int K, int N,
                              void* stream) {
    PGK_REQUIRE(M >= 1 && K >= 1 && N >= 1, "w8a16_gemm failed: invalid shape M=%d K=%d N=%d", M, K, N);
    constexpr int MT = 8;
    dim3 grid((N + 127) / 128, (M + MT - 1) / MT);
    PGK_CUDA(pgk::launch(pgk::w8a16_gemm_kn_kernel<MT>, grid, dim3(pgk::kGW * 32), 0, (cudaStream_t)stream,
                         (const uint16_t*)A, (const uint8_t*)B_kn, (const uint16_t*)S, (uint16_t*)C, M, K, N));
    return 0;
}
