/*
 * xattn_fusion_debug.h — TEST-ONLY entry points of libxattn_fusion_debug.so (the product library libxattn_fusion.so plus
 * these): single kernels of the product in isolation and the tcgen05 hardware probe the GPU tests pin.  Nothing here is on
 * the product path and none of it replaces reference code.
 */
#ifndef XATTN_FUSION_DEBUG_H_
#define XATTN_FUSION_DEBUG_H_

#include "xattn_fusion.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One tensor-core GEMM of the training path in isolation (tests): form 0: C[M,N] = A[M,K]·W[N,K]^T + bias; form 1: C[M,N] =
 * A[M,K]·W[K,N]; form 2: C[N1=K... see gemm_tc.cuh] C[N,K] += A[M,N]^T·B[M,K] with fp32 operands; form 3: the same with bf16
 * operands.  C is fp32. */
XATTN_API int xattn_debug_gemm_tf32(int32_t form, const void* A, const void* B, const float* bias, float* C, int32_t M, int32_t N,
                                    int32_t K, int32_t relu, int32_t accumulate, void* stream);

/* The tcgen05/TMA projection GEMM on its own: Y[M,N] = X[M,K]·W[N,K]^T + bias (X, W bf16; Y bf16 or fp32 with row
 * pitch out_pitch in elements; N % 128 == 0, K % 64 == 0).  This is the kernel behind v_in_proj (fusion.py:372) and
 * a_in_proj∘audio_seq_proj (:378, :386) in xattn_forward(precision = BF16); exported so tests can pin it in isolation. */
XATTN_API int xattn_debug_gemm_tc(const void* X, const void* W, const float* bias, void* Y, int32_t out_bf16, int64_t out_pitch,
                        int32_t M, int32_t N, int32_t K, void* stream);

/* Hardware probe of the tcgen05 features the attention kernels use (M = 64 accumulators in both halves of a TMEM
 * sub-partition, MN-major B operand, hand-swizzled operand tiles, 16x256b TMEM loads): two products in one CTA,
 * D_lo = A[0:64]·op(B[0]), D_hi = A[64:128]·op(B[1]); A [128,K] bf16; mode 0: B [2,N,K] (K-major), mode 1: B [2,K,N]
 * (MN-major).  out32 [128,N] = raw TMEM lanes x columns (32x32b loads), out16 [2,4,32,64] = raw 16x256b.x16 registers
 * (N >= 128).  Test-only; not on the product path. */
XATTN_API int xattn_debug_umma_probe(const void* A, const void* B, float* out32, float* out16, int32_t N, int32_t K,
                           int32_t mode, void* stream);

/* The dropout factors (0 or 1/keep) the training kernels draw for Philox stream `rng_stream` (1 attention v2a, 2 attention a2v,
 * 3 / 4 stochastic depth video / audio, 5 head, 6 prior_net, 7 / 8 attention pooler video / audio) and element indices
 * [idx0, idx0 + n); seed / offset as in xattn_config.  out: n floats. */
XATTN_API int xattn_debug_dropout_factors(float p, uint64_t seed, uint64_t offset, uint32_t rng_stream, uint64_t idx0, int64_t n,
                                          float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* XATTN_FUSION_DEBUG_H_ */
