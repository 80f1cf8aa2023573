/* oevae_test.h -- entry points of liboevae.so that exist for the kernel-level tests and the tuning scripts under bench/,
 * NOT part of the drop-in boundary (include/oevae.h).  Same conventions as oevae.h. */
#ifndef OEVAE_TEST_H
#define OEVAE_TEST_H
#include "oevae.h"
#ifdef __cplusplus
extern "C" {
#endif

/* The layered Dense contraction on its own (kernel-level tests): C[R,Cn] = sum_t A[r*sAr+t*sAt] * B[t*sBt+c*sBc]
 * (+bias[c], ReLU) in the handle's precision.  kind 0: forward (sAt==1, sBc==1), 1: dX (sAt==1, sBt==1),
 * 2: dW (sAr==1, sBc==1) -- the three operand orientations of models.py:31-34 and its tape gradient. */
OEVAE_API int oevae_gemm(oevae_t h, int32_t kind, const float* A, int64_t sAr, int64_t sAt, const float* B, int64_t sBt, int64_t sBc,
               float* C, int32_t ldc, int32_t R, int32_t Cn, int32_t T, const float* bias, int32_t relu, void* stream);

/* Hardware-layout probe (tests): C = A^T B through MN-major SWIZZLE_128B descriptors, A [K,M], B [K,N]; dumps the
 * raw TMEM accumulator [128 lanes, N] so that the M=64 / M=128 lane mapping can be asserted. */
OEVAE_API int oevae_probe_mn(const float* A, const float* B, int32_t K, int32_t M, int32_t N, float* raw_out, int32_t device, void* stream);

/* Tuning probe: cycles from issuing nm = {1,3,6,12,18,36} dependent tcgen05.mma (M x N x 16, bf16) to the mbarrier
 * wake-up after tcgen05.commit; out_host6 is a HOST array of 6 int64. */
OEVAE_API int oevae_probe_latency(int32_t M, int32_t N, int64_t* out_host6, int32_t device);

/* Tuning probe: cycles for nw warps to each issue nm tcgen05.mma (mode 0: descriptors rebuilt, 1: hoisted) and see
 * their commit; out_host1 is a HOST int64. */
OEVAE_API int oevae_probe_issue(int32_t M, int32_t N, int32_t nw, int32_t nm, int32_t mode, int64_t* out_host1, int32_t device);

/* Tuning probe: cycles for `depth` cp.async.bulk global->shared copies of `bytes` each, issued back to back by one
 * thread of each of n_cta CTAs (same_src: every CTA reads the same bytes, as the weight images are read);
 * out_host2 = {best of 5, first (cold)} of CTA 0. */
OEVAE_API int oevae_probe_bulk(int32_t bytes, int32_t depth, int32_t n_cta, int32_t same_src, int64_t* out_host2, int32_t device);

/* Phase stamps of the last oevae_train_step_dp update kernel (handles created with OEVAE_DP_STAMPS=1): out_host int64 [n_blocks][8] =
 * globaltimer ns at {entry, dependency met, local sum ready, sums pushed, all peer sums seen, exit, unused, unused}; returns n_blocks. */
OEVAE_API int oevae_dp_stamps_host(oevae_t h, int64_t* out_host, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif
