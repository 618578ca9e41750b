/*
 * oracle.c -- CPU restatement of PyGPUkit's decode hot path (TEST INFRASTRUCTURE).
 *
 * This file is the *checker*, never the product: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The shipped path
 * (pygpukit_b200/) never imports, links or executes anything under oracle/.
 *
 * Every function restates one reference CUDA kernel in plain C, keeping the
 * reference's accumulation ORDER (lane-strided partial sums + shfl_down tree, serial-K,
 * ...) so that results are deterministic and as close to the reference kernels as a
 * CPU can get.  File:line citations are relative to /root/reference.
 *
 * Parity pin: (1) tests/golden/*.npz generated from the reference's own Python CPU
 * functions (llm/quant.py, ops/nn/*.py _cpu branches, llm/layers/rope.py,
 * llm/sampling.py) by tests/golden/make_golden.py; (2) the reference's CUDA kernels
 * themselves, compiled in place by oracle/build_ref.sh into oracle/_ref/libref_cuda.so
 * and run on the B200 next to this oracle in tests/test_gpu_gemv.py / tests/test_gpu_nn.py.
 *
 * Known, documented deviations from the reference build (native/CMakeLists.txt:122
 * uses --use_fast_math): expf / division here are IEEE; the CUDA reference uses
 * ex2.approx / rcp.approx.  They differ in the last 1-2 ulp of f32, far below the
 * BF16 output rounding; the NVF4 quantiser can flip a nibble only when a value sits
 * within 2 ulp of a decision threshold.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------- */
/* scalar helpers                                                              */
/* ------------------------------------------------------------------------- */

static inline float bf16_to_f32(uint16_t h) {
    uint32_t u = ((uint32_t)h) << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

/* round-to-nearest-even, NaN kept quiet: same as __float2bfloat16_rn and the host
 * bit trick in core/array.py:386-395 for all finite inputs. */
static inline uint16_t f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

/* round a double straight to bf16 (single rounding; used for native bf16 add/mul). */
static inline uint16_t f64_to_bf16(double d) {
    float f = (float)d; /* RNE to f32 */
    /* fix double rounding: if f32 rounding landed exactly on a bf16 tie, re-decide
     * using the sign of the residual. */
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0xffffu) == 0x8000u && isfinite(f)) {
        double r = d - (double)f;
        if (r != 0.0) {
            /* true value is slightly above/below the tie */
            int up = ((r > 0) == (f > 0));
            uint16_t h = (uint16_t)(u >> 16);
            return up ? (uint16_t)(h + 1) : h;
        }
    }
    return f32_to_bf16(f);
}

static inline float f16_to_f32(uint16_t h) {
    uint32_t s = (h >> 15) & 1, e = (h >> 10) & 0x1f, m = h & 0x3ff, u;
    if (e == 0) {
        if (m == 0) u = s << 31;
        else {
            int sh = 0;
            while (!(m & 0x400)) { m <<= 1; sh++; }
            m &= 0x3ff;
            u = (s << 31) | ((uint32_t)(127 - 15 - sh + 1) << 23) | (m << 13);
        }
    } else if (e == 31) u = (s << 31) | 0x7f800000u | (m << 13);
    else u = (s << 31) | ((e + 112) << 23) | (m << 13);
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static inline uint16_t f32_to_f16(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    int32_t e = (int32_t)((x >> 23) & 0xff) - 127 + 15;
    uint32_t m = x & 0x7fffffu;
    if (((x >> 23) & 0xff) == 0xff) return (uint16_t)(sign | 0x7c00u | (m ? 0x200u : 0));
    if (e >= 31) return (uint16_t)(sign | 0x7c00u);
    if (e <= 0) {
        if (e < -10) return (uint16_t)sign;
        m |= 0x800000u;
        uint32_t shift = (uint32_t)(14 - e);
        uint32_t hm = m >> shift;
        uint32_t rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (hm & 1))) hm++;
        return (uint16_t)(sign | hm);
    }
    uint32_t hm = m >> 13, rem = m & 0x1fffu;
    uint16_t h = (uint16_t)(sign | ((uint32_t)e << 10) | hm);
    if (rem > 0x1000u || (rem == 0x1000u && (hm & 1))) h++;
    return h;
}

/* FP8 E4M3 decode as executed by every reference CUDA LUT:
 * native/ops/matmul/gemv/w8a16_bf16/sm120/fp8.cuh:57-122 (0x7F -> +480, 0xFF -> -480),
 * runtime LUTs w8a16_gemm.cu:384-400, grouped_gemm.cu:149-165. */
static inline float fp8_e4m3_decode(uint8_t v) {
    int s = v >> 7, e = (v >> 3) & 0xf, m = v & 7;
    float r = (e == 0) ? ldexpf((float)m, -9) : ldexpf(1.0f + (float)m / 8.0f, e - 7);
    return s ? -r : r;
}

/* UE4M3 scale decode: nvf4.cuh:65-115 (exp==0 treated as normal, bit 7 ignored). */
static inline float ue4m3_decode(uint8_t v) {
    int e = (v >> 3) & 0xf, m = v & 7;
    return ldexpf(1.0f + (float)m / 8.0f, e - 7);
}

/* E2M1 table: nvf4.cuh:40-43 */
static const float E2M1[16] = {0.0f, 0.5f, 1.0f, 1.5f, 2.0f, 3.0f, 4.0f, 6.0f,
                               0.0f, -0.5f, -1.0f, -1.5f, -2.0f, -3.0f, -4.0f, -6.0f};

/* warp shfl_down tree on a 32-lane register image; lane 0 carries the result.
 * (fp8_opt.cuh:203-207, bf16_opt.cuh:222-226, norm_kernels.cuh:553-555) */
static inline float warp_reduce_down(float v[32]) {
    for (int off = 16; off > 0; off >>= 1) {
        float t[32];
        for (int l = 0; l < 32; ++l) t[l] = v[l] + ((l + off < 32) ? v[l + off] : v[l]);
        memcpy(v, t, sizeof(t));
    }
    return v[0];
}

static inline float warp_max_down(float v[32]) {
    for (int off = 16; off > 0; off >>= 1) {
        float t[32];
        for (int l = 0; l < 32; ++l) t[l] = fmaxf(v[l], (l + off < 32) ? v[l + off] : v[l]);
        memcpy(v, t, sizeof(t));
    }
    return v[0];
}

void orc_fp8_e4m3_table(float* out256) {
    for (int i = 0; i < 256; ++i) out256[i] = fp8_e4m3_decode((uint8_t)i);
}
void orc_ue4m3_table(float* out256) {
    for (int i = 0; i < 256; ++i) out256[i] = ue4m3_decode((uint8_t)i);
}
void orc_e2m1_table(float* out16) { memcpy(out16, E2M1, sizeof(E2M1)); }

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* bench.py's CPU legs set the team size themselves: torchrun exports OMP_NUM_THREADS=1 to every rank */
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n >= 1) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ------------------------------------------------------------------------- */
/* W8A16 FP8 GEMV  (a-2, a-3)                                                  */
/* ------------------------------------------------------------------------- */

/* One output row, reference order.
 * K >= 128: gemv_fp8_warp_reduce_vec4_kernel fp8_opt.cuh:133-212 -- lane l owns
 *   k = 128*i + 4*l + {0..3}, tail (K%4) strided by 32, one f32 accumulator per lane,
 *   acc = fmaf(a, LUT[b]*scale, acc), then the shfl_down tree.
 * K < 128: gemv_fp8_warp_reduce_kernel fp8_opt.cuh:63-124 -- lane l owns k = l + 32*i.
 * (launcher picks by K: fp8_opt_kernels.cu:33-41). */
static float fp8_row_dot(const uint16_t* A, const uint8_t* Brow, const uint16_t* Srow,
                         int K, const float* lut) {
    float lane[32];
    for (int l = 0; l < 32; ++l) lane[l] = 0.0f;
    if (K >= 128) {
        int K4 = K & ~3;
        for (int l = 0; l < 32; ++l) {
            float acc = 0.0f;
            for (int kb = l * 4; kb < K4; kb += 128) {
                float scale = bf16_to_f32(Srow[kb / 128]);
                for (int j = 0; j < 4; ++j)
                    acc = fmaf(bf16_to_f32(A[kb + j]), lut[Brow[kb + j]] * scale, acc);
            }
            for (int k = K4 + l; k < K; k += 32) {
                float scale = bf16_to_f32(Srow[k / 128]);
                acc = fmaf(bf16_to_f32(A[k]), lut[Brow[k]] * scale, acc);
            }
            lane[l] = acc;
        }
    } else {
        for (int l = 0; l < 32; ++l) {
            float acc = 0.0f;
            for (int k = l; k < K; k += 32) {
                float scale = bf16_to_f32(Srow[k / 128]);
                acc = fmaf(bf16_to_f32(A[k]), lut[Brow[k]] * scale, acc);
            }
            lane[l] = acc;
        }
    }
    return warp_reduce_down(lane);
}

/* C[M,N] = A[M,K] . dequant(B[N,K])^T ; S is [ceil(N/128), ceil(K/128)] bf16.
 * M==1 is ops.gemv_fp8_bf16 (ops/matmul/gemv.py:73-131), M>1 is
 * ops.gemv_fp8_bf16_batched (:134-195; kernel fp8_opt.cuh:220-295 = same per-row order). */
void orc_gemv_fp8_bf16(const uint16_t* A, const uint8_t* B, const uint16_t* S, uint16_t* C,
                       int M, int K, int N) {
    float lut[256];
    orc_fp8_e4m3_table(lut);
    int sk = (K + 127) / 128;
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n) {
        const uint16_t* Srow = S + (size_t)(n / 128) * sk;
        for (int m = 0; m < M; ++m)
            C[(size_t)m * N + n] =
                f32_to_bf16(fp8_row_dot(A + (size_t)m * K, B + (size_t)n * K, Srow, K, lut));
    }
}

/* Block dequant: W[N,K] f32 = LUT[B] * S[n/128,k/128]  (llm/quant.py:323-367 and
 * layers/linear.py:181-211 do the same on the host with NaN at 0x7F/0xFF; the CUDA
 * value +-480 is used here, see header). */
void orc_dequant_fp8_block(const uint8_t* B, const uint16_t* S, float* W, int N, int K) {
    float lut[256];
    orc_fp8_e4m3_table(lut);
    int sk = (K + 127) / 128;
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k)
            W[(size_t)n * K + k] =
                lut[B[(size_t)n * K + k]] * bf16_to_f32(S[(size_t)(n / 128) * sk + k / 128]);
}

/* ------------------------------------------------------------------------- */
/* W8A16 GEMM, B[K,N] layout (a-4)                                             */
/* ------------------------------------------------------------------------- */

/* ops.w8a16_gemm_sm120 (ops/matmul/w8a16.py:38-119) -> pygpukit_w8a16_gemm_sm120
 * (w8a16_gemm.cu:417-456).
 *  M < 16 or K < 32 : w8a16_gemm_scalar_kernel (:341-367): acc += a*b serial over k,
 *                     b = LUT[B[k,n]] * S[k/128, n/128]; nvcc contracts to fma.
 *  otherwise        : w8a16_gemm_kernel_bf16tc (:104-335): the dequantised weight is
 *                     ROUNDED TO BF16 (:200-203) and fed to mma.sync m16n8k16 with f32
 *                     accumulate.  Products bf16*bf16 are exact in f32; the in-MMA
 *                     summation order is hardware-defined, so this path is restated as
 *                     a plain k-ordered f32 sum and compared with a tolerance. */
void orc_w8a16_gemm(const uint16_t* A, const uint8_t* B, const uint16_t* S, uint16_t* C,
                    int M, int N, int K) {
    float lut[256];
    orc_fp8_e4m3_table(lut);
    int sn = (N + 127) / 128;
    int tc = !(M < 16 || K < 32);
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n) {
        for (int m = 0; m < M; ++m) {
            float acc = 0.0f;
            for (int k = 0; k < K; ++k) {
                float a = bf16_to_f32(A[(size_t)m * K + k]);
                float b = lut[B[(size_t)k * N + n]] *
                          bf16_to_f32(S[(size_t)(k / 128) * sn + n / 128]);
                if (tc) b = bf16_to_f32(f32_to_bf16(b));
                acc = fmaf(a, b, acc);
            }
            C[(size_t)m * N + n] = f32_to_bf16(acc);
        }
    }
}

/* grouped GEMM for MoE experts: grouped_gemm_tiled_kernel grouped_gemm.cu:77-142.
 * row m uses expert row_expert_ids[m] of B[E,N,K] / S[E,ceil(N/128),ceil(K/128)];
 * acc += a*b serial over k (contracted to fma). */
void orc_grouped_gemm_fp8_bf16(const uint16_t* A, const uint8_t* B, const uint16_t* S,
                               uint16_t* C, const int32_t* row_expert_ids, int M, int N,
                               int K) {
    float lut[256];
    orc_fp8_e4m3_table(lut);
    int sn = (N + 127) / 128, sk = (K + 127) / 128;
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n) {
        for (int m = 0; m < M; ++m) {
            int e = row_expert_ids[m];
            const uint8_t* Be = B + (size_t)e * N * K;
            const uint16_t* Se = S + (size_t)e * sn * sk;
            float acc = 0.0f;
            for (int k = 0; k < K; ++k) {
                float b = lut[Be[(size_t)n * K + k]] *
                          bf16_to_f32(Se[(size_t)(n / 128) * sk + k / 128]);
                acc = fmaf(bf16_to_f32(A[(size_t)m * K + k]), b, acc);
            }
            C[(size_t)m * N + n] = f32_to_bf16(acc);
        }
    }
}

/* ------------------------------------------------------------------------- */
/* BF16 GEMV (a-16)                                                            */
/* ------------------------------------------------------------------------- */

/* launch_gemv_bf16_opt bf16_opt.cuh (K>=4096 -> vec8 kernel :134-231, else :56-127). */
static float bf16_row_dot(const uint16_t* A, const uint16_t* Brow, int K) {
    float lane[32];
    if (K >= 4096) {
        int Kloop = K & ~255, K4 = K & ~3;
        for (int l = 0; l < 32; ++l) {
            float a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            for (int kb = l * 8; kb < Kloop; kb += 256) {
                a0 = fmaf(bf16_to_f32(A[kb + 0]), bf16_to_f32(Brow[kb + 0]), a0);
                a0 = fmaf(bf16_to_f32(A[kb + 1]), bf16_to_f32(Brow[kb + 1]), a0);
                a1 = fmaf(bf16_to_f32(A[kb + 2]), bf16_to_f32(Brow[kb + 2]), a1);
                a1 = fmaf(bf16_to_f32(A[kb + 3]), bf16_to_f32(Brow[kb + 3]), a1);
                a2 = fmaf(bf16_to_f32(A[kb + 4]), bf16_to_f32(Brow[kb + 4]), a2);
                a2 = fmaf(bf16_to_f32(A[kb + 5]), bf16_to_f32(Brow[kb + 5]), a2);
                a3 = fmaf(bf16_to_f32(A[kb + 6]), bf16_to_f32(Brow[kb + 6]), a3);
                a3 = fmaf(bf16_to_f32(A[kb + 7]), bf16_to_f32(Brow[kb + 7]), a3);
            }
            for (int kb = Kloop + l * 4; kb < K4; kb += 128)
                for (int j = 0; j < 4; ++j)
                    a0 = fmaf(bf16_to_f32(A[kb + j]), bf16_to_f32(Brow[kb + j]), a0);
            for (int k = K4 + l; k < K; k += 32)
                a0 = fmaf(bf16_to_f32(A[k]), bf16_to_f32(Brow[k]), a0);
            lane[l] = a0 + a1 + a2 + a3;
        }
    } else {
        int Kloop = K & ~127;
        for (int l = 0; l < 32; ++l) {
            float acc = 0.0f;
            for (int kb = l * 4; kb < Kloop; kb += 128)
                for (int j = 0; j < 4; ++j)
                    acc = fmaf(bf16_to_f32(A[kb + j]), bf16_to_f32(Brow[kb + j]), acc);
            for (int k = Kloop + l; k < K; k += 32)
                acc = fmaf(bf16_to_f32(A[k]), bf16_to_f32(Brow[k]), acc);
            lane[l] = acc;
        }
    }
    return warp_reduce_down(lane);
}

void orc_gemv_bf16(const uint16_t* A, const uint16_t* B, uint16_t* C, int M, int K, int N) {
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n)
        for (int m = 0; m < M; ++m)
            C[(size_t)m * N + n] =
                f32_to_bf16(bf16_row_dot(A + (size_t)m * K, B + (size_t)n * K, K));
}

/* ------------------------------------------------------------------------- */
/* NVF4 (E2M1 x UE4M3, block 32) GEMV + quantiser (a-5, a-6)                   */
/* ------------------------------------------------------------------------- */

/* ops.gemv_nvf4_bf16 (ops/matmul/nvf4.py:132-194) -> pygpukit_gemv_nvf4_bf16
 * (nvf4.cu:161-178) -> launch_gemv_nvf4_bf16 (nvf4_kernels.cu:206-233).
 * Thread n walks k = 0..K-1 in order; per 32-block lut[i] = E2M1[i]*scale in f32
 * (index 0 and 8 are +0.0f: nvf4_kernels.cu:109,117); low nibble = even k;
 * acc = fmaf(a, lut[nib], acc); output bf16(alpha*acc) (:79-200).  The generic
 * kernel (:19-73, K%32 != 0) uses E2M1[i]*scale per element -- same values. */
void orc_gemv_nvf4_bf16(const uint16_t* A, const uint8_t* Bd, const uint8_t* Bs, uint16_t* C,
                        int K, int N, float alpha) {
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n) {
        float acc = 0.0f;
        int nsb = (K + 31) / 32;
        for (int sb = 0; sb < nsb; ++sb) {
            float scale = ue4m3_decode(Bs[(size_t)sb * N + n]);
            float lut[16];
            for (int i = 0; i < 16; ++i) lut[i] = E2M1[i] * scale;
            lut[0] = 0.0f;
            lut[8] = 0.0f; /* +0 in both kernels (LUT[8] is +0.0f) */
            int k0 = sb * 32, k1 = k0 + 32 < K ? k0 + 32 : K;
            for (int k = k0; k < k1; k += 2) {
                uint8_t p = Bd[(size_t)(k / 2) * N + n];
                float a0 = bf16_to_f32(A[k]);
                float a1 = (k + 1 < K) ? bf16_to_f32(A[k + 1]) : 0.0f;
                acc = fmaf(a0, lut[p & 0xf], acc);
                acc = fmaf(a1, lut[(p >> 4) & 0xf], acc);
            }
        }
        C[n] = f32_to_bf16(alpha * acc);
    }
}

/* Dequantise NVF4 to f32 [K,N] (exact: 2-bit x 4-bit significands). */
void orc_dequant_nvf4(const uint8_t* Bd, const uint8_t* Bs, float* W, int K, int N) {
#pragma omp parallel for schedule(static)
    for (int k = 0; k < K; ++k)
        for (int n = 0; n < N; ++n) {
            uint8_t p = Bd[(size_t)(k / 2) * N + n];
            int nib = (k & 1) ? (p >> 4) & 0xf : p & 0xf;
            W[(size_t)k * N + n] = E2M1[nib] * ue4m3_decode(Bs[(size_t)(k / 32) * N + n]);
        }
}

static inline uint8_t nvf4_snap(float v) { /* nvf4_kernels.cu:303-314 */
    uint8_t s = (v < 0) ? 0x8 : 0x0;
    v = fabsf(v);
    if (v < 0.25f) return s | 0;
    if (v < 0.75f) return s | 1;
    if (v < 1.25f) return s | 2;
    if (v < 1.75f) return s | 3;
    if (v < 2.5f) return s | 4;
    if (v < 3.5f) return s | 5;
    if (v < 5.0f) return s | 6;
    return s | 7;
}

/* ops.quantize_bf16_to_nvf4 (ops/matmul/nvf4.py:32-77) -> quantize_bf16_to_nvf4_kernel
 * (nvf4_kernels.cu:239-324).  input [K,N] bf16 row-major -> data [K/2,N], scale
 * [ceil(K/32),N].  rowmajor != 0 restates the _rowmajor kernel (:357-446): data [N,K/2],
 * scale [N,ceil(K/32)].
 * The reference build uses --use_fast_math: max_abs / 6.0f and 1.0f / encoded_scale
 * become approximate; encoded_scale is a 4-bit-significand power-of-two multiple, its
 * reciprocal and the products are compared against thresholds, so a flip needs a value
 * within 2 ulp of a threshold (see header). */
void orc_quantize_bf16_to_nvf4(const uint16_t* in, uint8_t* data, uint8_t* scale_out, int K,
                               int N, int rowmajor) {
    int nsb = (K + 31) / 32;
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; ++n) {
        for (int sb = 0; sb < nsb; ++sb) {
            int k0 = sb * 32, k1 = k0 + 32 < K ? k0 + 32 : K;
            float max_abs = 0.0f;
            for (int k = k0; k < k1; ++k)
                max_abs = fmaxf(max_abs, fabsf(bf16_to_f32(in[(size_t)k * N + n])));
            float scale = (max_abs > 1e-8f) ? (max_abs / 6.0f) : 1.0f;
            int exp_raw = 0;
            float norm = scale;
            if (norm >= 2.0f) {
                while (norm >= 2.0f && exp_raw < 8) { norm *= 0.5f; exp_raw++; }
            } else if (norm < 1.0f && norm > 1e-8f) {
                while (norm < 1.0f && exp_raw > -7) { norm *= 2.0f; exp_raw--; }
            }
            int mant = (int)rintf((norm - 1.0f) * 8.0f);
            if (mant < 0) mant = 0;
            if (mant > 7) mant = 7;
            int eb = exp_raw + 7;
            if (eb < 0) eb = 0;
            if (eb > 15) eb = 15;
            uint8_t enc = (uint8_t)(((eb & 0xf) << 3) | (mant & 7));
            if (rowmajor) scale_out[(size_t)n * nsb + sb] = enc;
            else scale_out[(size_t)sb * N + n] = enc;
            float es = (1.0f + (float)mant / 8.0f) * ldexpf(1.0f, eb - 7);
            float inv = 1.0f / es;
            for (int k = k0; k < k1; k += 2) {
                float v0 = bf16_to_f32(in[(size_t)k * N + n]) * inv;
                float v1 = (k + 1 < k1) ? bf16_to_f32(in[(size_t)(k + 1) * N + n]) * inv : 0.0f;
                uint8_t b = (uint8_t)((nvf4_snap(v1) << 4) | (nvf4_snap(v0) & 0xf));
                if (rowmajor) data[(size_t)n * (K / 2) + k / 2] = b;
                else data[(size_t)(k / 2) * N + n] = b;
            }
        }
    }
}

/* ------------------------------------------------------------------------- */
/* block reductions in the reference's order                                   */
/* ------------------------------------------------------------------------- */

/* "thread-strided partials -> per-warp shfl_down -> smem[warp] -> warp 0 shfl_down"
 * (norm_kernels.cuh:545-570, attention_kernels.cuh:306-352).  part[] holds one partial
 * per thread (nthreads <= 1024, multiple of 32 in every reference launch). */
static float block_reduce_sum(const float* part, int nthreads) {
    float warp_tot[32];
    int nw = (nthreads + 31) / 32;
    for (int w = 0; w < 32; ++w) warp_tot[w] = 0.0f;
    for (int w = 0; w < nw; ++w) {
        float v[32];
        for (int l = 0; l < 32; ++l) v[l] = (w * 32 + l < nthreads) ? part[w * 32 + l] : 0.0f;
        warp_tot[w] = warp_reduce_down(v);
    }
    return warp_reduce_down(warp_tot);
}

static float block_reduce_max(const float* part, int nthreads) {
    float warp_tot[32];
    int nw = (nthreads + 31) / 32;
    for (int w = 0; w < 32; ++w) warp_tot[w] = -INFINITY;
    for (int w = 0; w < nw; ++w) {
        float v[32];
        for (int l = 0; l < 32; ++l)
            v[l] = (w * 32 + l < nthreads) ? part[w * 32 + l] : -INFINITY;
        warp_tot[w] = warp_max_down(v);
    }
    return warp_max_down(warp_tot);
}

static int norm_block_size(int features) { /* rmsnorm.inl:18-19, layernorm.inl same */
    int b = (features + 31) / 32 * 32;
    if (b > 256) b = 256;
    if (b < 32) b = 32;
    return b;
}

/* ------------------------------------------------------------------------- */
/* RMSNorm / LayerNorm (a-8)                                                   */
/* ------------------------------------------------------------------------- */

/* ops.rmsnorm (ops/nn/norm.py:118-218) -> rmsnorm_bf16_kernel norm_kernels.cuh:535-585.
 * f32: sum_sq += val*val (contracted to fma by nvcc), inv = rsqrtf(sum/features + eps),
 * y = bf16(x*inv*g) left to right. */
void orc_rmsnorm_bf16(const uint16_t* x, const uint16_t* gamma, uint16_t* y, int rows,
                      int features, float eps) {
    int bs = norm_block_size(features);
#pragma omp parallel for schedule(static)
    for (int r = 0; r < rows; ++r) {
        const uint16_t* xr = x + (size_t)r * features;
        float part[256];
        for (int t = 0; t < bs; ++t) {
            float s = 0.0f;
            for (int i = t; i < features; i += bs) {
                float v = bf16_to_f32(xr[i]);
                s = fmaf(v, v, s);
            }
            part[t] = s;
        }
        float sum = block_reduce_sum(part, bs);
        float inv = 1.0f / sqrtf(sum / (float)features + eps);
        for (int i = 0; i < features; ++i)
            y[(size_t)r * features + i] =
                f32_to_bf16(bf16_to_f32(xr[i]) * inv * bf16_to_f32(gamma[i]));
    }
}

/* rmsnorm_f32_kernel norm_kernels.cuh:375-425 (same structure, f32 I/O). */
void orc_rmsnorm_f32(const float* x, const float* gamma, float* y, int rows, int features,
                     float eps) {
    int bs = norm_block_size(features);
#pragma omp parallel for schedule(static)
    for (int r = 0; r < rows; ++r) {
        const float* xr = x + (size_t)r * features;
        float part[256];
        for (int t = 0; t < bs; ++t) {
            float s = 0.0f;
            for (int i = t; i < features; i += bs) s = fmaf(xr[i], xr[i], s);
            part[t] = s;
        }
        float sum = block_reduce_sum(part, bs);
        float inv = 1.0f / sqrtf(sum / (float)features + eps);
        for (int i = 0; i < features; ++i) y[(size_t)r * features + i] = xr[i] * inv * gamma[i];
    }
}

/* layernorm_f32_kernel norm_kernels.cuh:32-120: mean, then var = mean((x-mean)^2),
 * y = (x-mean)*rsqrt(var+eps)*g + b.  (GPT-2, config[0]) */
void orc_layernorm_f32(const float* x, const float* gamma, const float* beta, float* y,
                       int rows, int features, float eps) {
    int bs = norm_block_size(features);
#pragma omp parallel for schedule(static)
    for (int r = 0; r < rows; ++r) {
        const float* xr = x + (size_t)r * features;
        float part[256];
        for (int t = 0; t < bs; ++t) {
            float s = 0.0f;
            for (int i = t; i < features; i += bs) s += xr[i];
            part[t] = s;
        }
        float mean = block_reduce_sum(part, bs) / (float)features;
        for (int t = 0; t < bs; ++t) {
            float s = 0.0f;
            for (int i = t; i < features; i += bs) {
                float d = xr[i] - mean;
                s = fmaf(d, d, s);
            }
            part[t] = s;
        }
        float var = block_reduce_sum(part, bs) / (float)features;
        float inv = 1.0f / sqrtf(var + eps);
        for (int i = 0; i < features; ++i)
            y[(size_t)r * features + i] = (xr[i] - mean) * inv * gamma[i] + beta[i];
    }
}

/* ------------------------------------------------------------------------- */
/* RoPE (a-9)                                                                  */
/* ------------------------------------------------------------------------- */

/* rope_bf16_f32table_kernel elementwise_kernels.cuh:253-307: NeoX half-split rotation,
 * f32 math, f32 tables [seq, head_dim] (first half read), bf16 store.
 * q0*c - q1*sn is contracted by nvcc (-fmad=true) to fma(q0, c, -(q1*sn)); same for the
 * second output: fma(q1, c, q0*sn). */
static void rope_rows_bf16(uint16_t* x, const float* cosf_, const float* sinf_, int seq,
                           int heads, int hd) {
    int half = hd / 2;
    for (int s = 0; s < seq; ++s)
        for (int h = 0; h < heads; ++h) {
            uint16_t* p = x + ((size_t)s * heads + h) * hd;
            for (int d = 0; d < half; ++d) {
                float x0 = bf16_to_f32(p[d]), x1 = bf16_to_f32(p[d + half]);
                float c = cosf_[(size_t)s * hd + d], sn = sinf_[(size_t)s * hd + d];
                p[d] = f32_to_bf16(fmaf(x0, c, -(x1 * sn)));
                p[d + half] = f32_to_bf16(fmaf(x1, c, x0 * sn));
            }
        }
}

void orc_rope_bf16_f32table(uint16_t* q, uint16_t* k, const float* cos_, const float* sin_,
                            int seq, int hq, int hk, int hd) {
    rope_rows_bf16(q, cos_, sin_, seq, hq, hd);
    rope_rows_bf16(k, cos_, sin_, seq, hk, hd);
}

/* rope_f32_kernel elementwise_kernels.cuh:82-131 (f32 q/k/tables). */
static void rope_rows_f32(float* x, const float* cosf_, const float* sinf_, int seq, int heads,
                          int hd) {
    int half = hd / 2;
    for (int s = 0; s < seq; ++s)
        for (int h = 0; h < heads; ++h) {
            float* p = x + ((size_t)s * heads + h) * hd;
            for (int d = 0; d < half; ++d) {
                float x0 = p[d], x1 = p[d + half];
                float c = cosf_[(size_t)s * hd + d], sn = sinf_[(size_t)s * hd + d];
                p[d] = fmaf(x0, c, -(x1 * sn));
                p[d + half] = fmaf(x1, c, x0 * sn);
            }
        }
}

void orc_rope_f32(float* q, float* k, const float* cos_, const float* sin_, int seq, int hq,
                  int hk, int hd) {
    rope_rows_f32(q, cos_, sin_, seq, hq, hd);
    rope_rows_f32(k, cos_, sin_, seq, hk, hd);
}

/* ------------------------------------------------------------------------- */
/* KV-cache update (a-10)                                                      */
/* ------------------------------------------------------------------------- */

/* kv_cache_update_gqa_*_kernel kv_cache_kernels.cuh:176-260: new_kv [1,Hkv,hd] ->
 * cache [H, max_seq, hd], each kv head replicated H/Hkv times.  elem = bytes/element. */
void orc_kv_cache_update_gqa(const void* new_kv, void* cache, int num_heads, int num_kv_heads,
                             int head_dim, int max_seq, int position, int elem) {
    int groups = num_heads / num_kv_heads;
    for (int h = 0; h < num_heads; ++h)
        memcpy((char*)cache + ((size_t)h * max_seq + position) * head_dim * elem,
               (const char*)new_kv + (size_t)(h / groups) * head_dim * elem,
               (size_t)head_dim * elem);
}

/* kv_cache_prefill_gqa_* kv_cache_kernels.cuh:330-421: new_kv [S,Hkv,hd] -> rows
 * start_pos..start_pos+S-1 of every replicated head. */
void orc_kv_cache_prefill_gqa(const void* new_kv, void* cache, int num_heads, int num_kv_heads,
                              int head_dim, int max_seq, int start_pos, int seq_len, int elem) {
    int groups = num_heads / num_kv_heads;
    for (int h = 0; h < num_heads; ++h)
        for (int s = 0; s < seq_len; ++s)
            memcpy((char*)cache + ((size_t)h * max_seq + start_pos + s) * head_dim * elem,
                   (const char*)new_kv + ((size_t)s * num_kv_heads + h / groups) * head_dim * elem,
                   (size_t)head_dim * elem);
}

/* ------------------------------------------------------------------------- */
/* SDPA over a fixed cache (a-7)                                               */
/* ------------------------------------------------------------------------- */

/* load one element of a 16/32-bit float array as f32; kind: 0=f32, 1=f16, 2=bf16 */
static inline float ld_as_f32(const void* p, size_t i, int kind) {
    if (kind == 0) return ((const float*)p)[i];
    if (kind == 1) return f16_to_f32(((const uint16_t*)p)[i]);
    return bf16_to_f32(((const uint16_t*)p)[i]);
}
static inline void st_from_f32(void* p, size_t i, int kind, float v) {
    if (kind == 0) ((float*)p)[i] = v;
    else if (kind == 1) ((uint16_t*)p)[i] = f32_to_f16(v);
    else ((uint16_t*)p)[i] = f32_to_bf16(v);
}

/* ops.sdpa_causal_fixed_cache (ops/nn/attention.py:159-192) -> sdpa_causal_dispatch
 * (sdpa_causal.inl:452-669) -> sdpa_causal_{f32,f16,bf16}_kernel
 * (attention_kernels.cuh:32-370), block = 128 threads, one block per (head, q_pos).
 *   score_j = (sum_d q_d*k_jd serial over d, fma-contracted) * scale ; -inf if masked
 *   max     = thread-strided partial max -> block reduce
 *   e_j     = expf(score_j - max); sum = thread-strided partial sums -> block reduce
 *   p_j     = e_j * (1/sum)          (normalised BEFORE the PV product)
 *   out_d   = sum_j p_j * v_jd serial over j (all kv_len rows, masked ones contribute 0)
 * Q [H,q_len,hd], K/V [H,kv_stride,hd], out [H,q_len,hd]; causal_offset = kv_len-q_len. */
void orc_sdpa_causal_fixed_cache(const void* Q, const void* K, const void* V, void* out,
                                 int n_heads, int q_len, int kv_len, int kv_stride,
                                 int head_dim, float scale, int kind) {
    const int BS = 128;
    if (scale <= 0.0f) scale = 1.0f / sqrtf((float)head_dim);
    int causal_offset = kv_len - q_len;
#pragma omp parallel for collapse(2) schedule(static)
    for (int h = 0; h < n_heads; ++h) {
        for (int qp = 0; qp < q_len; ++qp) {
            size_t qoff = ((size_t)h * q_len + qp) * head_dim;
            size_t kvoff = (size_t)h * kv_stride * head_dim;
            int max_attend = causal_offset + qp + 1;
            if (max_attend > kv_len) max_attend = kv_len;
            float* sc = (float*)malloc(sizeof(float) * (size_t)kv_len);
            float part[128];
            for (int t = 0; t < BS; ++t) part[t] = -INFINITY;
            for (int j = 0; j < kv_len; ++j) {
                float s;
                if (j < max_attend) {
                    s = 0.0f;
                    for (int d = 0; d < head_dim; ++d)
                        s = fmaf(ld_as_f32(Q, qoff + d, kind),
                                 ld_as_f32(K, kvoff + (size_t)j * head_dim + d, kind), s);
                    s *= scale;
                } else s = -INFINITY;
                sc[j] = s;
                if (s > part[j % BS]) part[j % BS] = s;
            }
            float mx = block_reduce_max(part, BS);
            for (int t = 0; t < BS; ++t) part[t] = 0.0f;
            for (int j = 0; j < kv_len; ++j) {
                float e = expf(sc[j] - mx);
                sc[j] = e;
                part[j % BS] += e;
            }
            float sum = block_reduce_sum(part, BS);
            float inv = 1.0f / sum;
            for (int j = 0; j < kv_len; ++j) sc[j] *= inv;
            for (int d = 0; d < head_dim; ++d) {
                float o = 0.0f;
                for (int j = 0; j < kv_len; ++j)
                    o = fmaf(sc[j], ld_as_f32(V, kvoff + (size_t)j * head_dim + d, kind), o);
                st_from_f32(out, qoff + d, kind, o);
            }
            free(sc);
        }
    }
}

/* flash_decoding_f16 flash_decoding.cuh:75-377 (q_len == 1, chunk = 256 positions,
 * block = 256 threads): per chunk c: m_c = max_j s_j, l_c = sum_j exp(s_j - m_c),
 * o_c = sum_j exp(s_j - m_c) v_j (serial over j, UN-normalised); phase 2: M = max_c m_c,
 * L = sum_c l_c*exp(m_c - M) (serial over c), out = (sum_c o_c*exp(m_c - M)) * (1/L).
 * The reference enables it for FP16 only (sdpa_causal.inl:479-514); `kind` lets the
 * same order be evaluated for bf16 inputs. */
void orc_flash_decoding(const void* Q, const void* K, const void* V, void* out, int n_heads,
                        int kv_len, int kv_stride, int head_dim, float scale, int kind) {
    const int CH = 256, BS = 256;
    if (scale <= 0.0f) scale = 1.0f / sqrtf((float)head_dim);
    int nch = (kv_len + CH - 1) / CH;
#pragma omp parallel for schedule(static)
    for (int h = 0; h < n_heads; ++h) {
        size_t qoff = (size_t)h * head_dim, kvoff = (size_t)h * kv_stride * head_dim;
        float* po = (float*)calloc((size_t)nch * head_dim, sizeof(float));
        float* pm = (float*)malloc(sizeof(float) * nch);
        float* pl = (float*)malloc(sizeof(float) * nch);
        for (int c = 0; c < nch; ++c) {
            int c0 = c * CH, c1 = c0 + CH < kv_len ? c0 + CH : kv_len, len = c1 - c0;
            float sc[256], part[256];
            for (int t = 0; t < BS; ++t) part[t] = -3.402823466e+38f;
            for (int j = 0; j < len; ++j) {
                float s = 0.0f;
                for (int d = 0; d < head_dim; ++d)
                    s = fmaf(ld_as_f32(Q, qoff + d, kind),
                             ld_as_f32(K, kvoff + (size_t)(c0 + j) * head_dim + d, kind), s);
                s *= scale;
                sc[j] = s;
                part[j] = fmaxf(part[j], s);
            }
            float mx = -3.402823466e+38f;
            { /* warp_reduce_max uses xor-free shfl_down; max is order independent */
                for (int t = 0; t < BS; ++t) mx = fmaxf(mx, part[t]);
            }
            for (int t = 0; t < BS; ++t) part[t] = 0.0f;
            for (int j = 0; j < len; ++j) {
                sc[j] = expf(sc[j] - mx);
                part[j] += sc[j];
            }
            pl[c] = block_reduce_sum(part, BS);
            pm[c] = mx;
            for (int j = 0; j < len; ++j)
                for (int d = 0; d < head_dim; ++d)
                    po[(size_t)c * head_dim + d] =
                        fmaf(sc[j], ld_as_f32(V, kvoff + (size_t)(c0 + j) * head_dim + d, kind),
                             po[(size_t)c * head_dim + d]);
        }
        float M = -3.402823466e+38f, L = 0.0f;
        for (int c = 0; c < nch; ++c) M = fmaxf(M, pm[c]);
        for (int c = 0; c < nch; ++c) L = fmaf(pl[c], expf(pm[c] - M), L);
        float invL = (L > 0.0f) ? 1.0f / L : 0.0f;
        for (int d = 0; d < head_dim; ++d) {
            float o = 0.0f;
            for (int c = 0; c < nch; ++c) o = fmaf(po[(size_t)c * head_dim + d], expf(pm[c] - M), o);
            st_from_f32(out, qoff + d, kind, o * invL);
        }
        free(po); free(pm); free(pl);
    }
}

/* ------------------------------------------------------------------------- */
/* elementwise glue (rounding points of Appendix A.1)                          */
/* ------------------------------------------------------------------------- */

/* silu_bf16_kernel activation_kernels.cuh:40-42,164-171: bf16(x/(1+expf(-x))). */
void orc_silu_bf16(const uint16_t* x, uint16_t* y, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        float v = bf16_to_f32(x[i]);
        y[i] = f32_to_bf16(v / (1.0f + expf(-v)));
    }
}

/* gelu_f32 activation_kernels.cuh:26-31 (tanh approximation). */
void orc_gelu_f32(const float* x, float* y, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        float v = x[i], v3 = v * v * v;
        y[i] = v * 0.5f * (1.0f + tanhf(0.7978845608f * (v + 0.044715f * v3)));
    }
}

/* add_inplace_bf16_kernel = __hadd (single rounding) elementwise_kernels.cuh:381-390 */
void orc_add_inplace_bf16(uint16_t* a, const uint16_t* b, size_t n) {
    for (size_t i = 0; i < n; ++i)
        a[i] = f64_to_bf16((double)bf16_to_f32(a[i]) + (double)bf16_to_f32(b[i]));
}

/* mul_inplace_bf16_kernel = __hmul elementwise_kernels.cuh:429-438 (product exact in f32) */
void orc_mul_inplace_bf16(uint16_t* a, const uint16_t* b, size_t n) {
    for (size_t i = 0; i < n; ++i) a[i] = f32_to_bf16(bf16_to_f32(a[i]) * bf16_to_f32(b[i]));
}

/* add_bf16_kernel: f32 add -> bf16 (native/ops/elementwise/elementwise_kernels.cuh:55-60);
 * bias_add_bf16_kernel: same per row (nn/elementwise_kernels.cuh:57-68). */
void orc_add_bf16(const uint16_t* a, const uint16_t* b, uint16_t* c, size_t n) {
    for (size_t i = 0; i < n; ++i) c[i] = f32_to_bf16(bf16_to_f32(a[i]) + bf16_to_f32(b[i]));
}
void orc_bias_add_inplace_bf16(uint16_t* x, const uint16_t* bias, int rows, int features) {
    for (int r = 0; r < rows; ++r)
        for (int i = 0; i < features; ++i) {
            size_t o = (size_t)r * features + i;
            x[o] = f32_to_bf16(bf16_to_f32(x[o]) + bf16_to_f32(bias[i]));
        }
}

/* embedding_lookup_* embedding_kernels.cuh:27-61: row copy. */
void orc_embedding_lookup(const void* table, void* out, int hidden, int token, int elem) {
    memcpy(out, (const char*)table + (size_t)token * hidden * elem, (size_t)hidden * elem);
}

/* greedy: llm/sampling.py:59-61 (np.argmax over f32 logits, first max wins) ==
 * sample_argmax_* sampling_kernels.cuh:55-200 (lowest index on ties). */
int orc_argmax_bf16(const uint16_t* x, int n) {
    int bi = 0;
    float bv = bf16_to_f32(x[0]);
    for (int i = 1; i < n; ++i) {
        float v = bf16_to_f32(x[i]);
        if (v > bv) { bv = v; bi = i; }
    }
    return bi;
}
int orc_argmax_f32(const float* x, int n) {
    int bi = 0;
    for (int i = 1; i < n; ++i) if (x[i] > x[bi]) bi = i;
    return bi;
}

/* ------------------------------------------------------------------------- */
/* MoE routing (a-14)                                                          */
/* ------------------------------------------------------------------------- */

/* topk_with_indices_bf16_kernel topk_kernels.cuh:111-147: k passes of strict-> argmax
 * (lowest index wins ties), used entries set to -1e30; values stored as bf16. */
void orc_moe_topk_bf16(const uint16_t* logits, uint16_t* values, int32_t* indices,
                       int num_tokens, int num_experts, int k) {
    for (int t = 0; t < num_tokens; ++t) {
        float loc[128];
        for (int i = 0; i < num_experts; ++i) loc[i] = bf16_to_f32(logits[(size_t)t * num_experts + i]);
        for (int j = 0; j < k; ++j) {
            float mv = -1e30f;
            int mi = 0;
            for (int i = 0; i < num_experts; ++i)
                if (loc[i] > mv) { mv = loc[i]; mi = i; }
            values[(size_t)t * k + j] = f32_to_bf16(mv);
            indices[(size_t)t * k + j] = mi;
            loc[mi] = -1e30f;
        }
    }
}

/* softmax_topk_bf16_kernel topk_kernels.cuh:222-256 (k <= 8), in place. */
void orc_moe_softmax_topk_bf16(uint16_t* values, int num_tokens, int k) {
    for (int t = 0; t < num_tokens; ++t) {
        float loc[8], mx, sum = 0.0f;
        for (int i = 0; i < k; ++i) loc[i] = bf16_to_f32(values[(size_t)t * k + i]);
        mx = loc[0];
        for (int i = 1; i < k; ++i) mx = fmaxf(mx, loc[i]);
        for (int i = 0; i < k; ++i) { loc[i] = expf(loc[i] - mx); sum += loc[i]; }
        float inv = 1.0f / sum;
        for (int i = 0; i < k; ++i) values[(size_t)t * k + i] = f32_to_bf16(loc[i] * inv);
    }
}

/* count_tokens_per_expert + compute_expert_offsets + build_permute_indices +
 * build_reverse_perm (permute_kernels.cuh:21-88,245-256).  The reference fills each
 * expert's segment with atomics (intra-expert order nondeterministic); this restatement
 * uses ascending flat index, one of the orders the reference can produce.  Everything
 * downstream (gather -> grouped GEMM -> scatter) is invariant to that order. */
void orc_moe_compute_permutation(const int32_t* expert_indices, int32_t* expert_counts,
                                 int32_t* expert_offsets, int32_t* permute_indices,
                                 int32_t* reverse_perm, int num_tokens, int num_experts, int k) {
    int total = num_tokens * k;
    for (int e = 0; e < num_experts; ++e) expert_counts[e] = 0;
    for (int i = 0; i < total; ++i) {
        int e = expert_indices[i];
        if (e >= 0 && e < num_experts) expert_counts[e]++;
    }
    int off = 0;
    for (int e = 0; e < num_experts; ++e) { expert_offsets[e] = off; off += expert_counts[e]; }
    expert_offsets[num_experts] = off;
    int32_t* wr = (int32_t*)calloc(num_experts, sizeof(int32_t));
    for (int i = 0; i < total; ++i) {
        int e = expert_indices[i];
        if (e >= 0 && e < num_experts) {
            int p = expert_offsets[e] + wr[e]++;
            permute_indices[p] = i;
            reverse_perm[i] = p;
        }
    }
    free(wr);
}

/* expand_expert_offsets_kernel moe_kernels.cuh:236-259 */
void orc_moe_expand_expert_offsets(const int32_t* expert_offsets, int32_t* row_expert_ids,
                                   int num_experts, int m_total) {
    for (int r = 0; r < m_total; ++r) {
        int lo = 0, hi = num_experts;
        while (lo < hi) {
            int mid = (lo + hi) / 2;
            if (expert_offsets[mid + 1] <= r) lo = mid + 1; else hi = mid;
        }
        row_expert_ids[r] = lo;
    }
}

/* gather_hidden_states_kernel permute_kernels.cuh:95-182: gathered[p] = hidden[perm[p]/k] */
void orc_moe_gather_bf16(const uint16_t* hidden, const int32_t* permute_indices,
                         uint16_t* gathered, int num_tokens, int hidden_size, int k) {
    for (int p = 0; p < num_tokens * k; ++p)
        memcpy(gathered + (size_t)p * hidden_size,
               hidden + (size_t)(permute_indices[p] / k) * hidden_size,
               (size_t)hidden_size * 2);
}

/* scatter_with_reverse_perm_kernel permute_kernels.cuh:218-242: f32 sum over slots 0..k-1
 * of weight*expert_out (contracted to fma), stored as bf16. */
void orc_moe_scatter_bf16(const uint16_t* expert_out, const uint16_t* router_weights,
                          const int32_t* reverse_perm, uint16_t* output, int num_tokens,
                          int hidden_size, int k) {
    for (int t = 0; t < num_tokens; ++t)
        for (int h = 0; h < hidden_size; ++h) {
            float sum = 0.0f;
            for (int s = 0; s < k; ++s) {
                int f = t * k + s;
                sum = fmaf(bf16_to_f32(router_weights[f]),
                           bf16_to_f32(expert_out[(size_t)reverse_perm[f] * hidden_size + h]), sum);
            }
            output[(size_t)t * hidden_size + h] = f32_to_bf16(sum);
        }
}
