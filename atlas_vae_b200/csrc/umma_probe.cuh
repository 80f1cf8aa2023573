// Probe kernel (tests only): validates the MN-major SWIZZLE_128B descriptors and the M=64 / M=128 TMEM accumulator
// layouts that the fused training kernel relies on.  C[M,N] = A^T B with A [K, M] (M contiguous), B [K, N]
// (N contiguous), both staged as [K rows x 64-column blocks] (the activation-tile layout) and consumed MN-major.
#pragma once
#include "umma.cuh"

namespace oevae {

// MN-major SWIZZLE_128B descriptor: rows = K (128 B each, 8-row groups SBO = 1024 B apart),
// 64-element MN blocks lbo_bytes apart.
__device__ __forceinline__ uint64_t smem_desc_sw128_mn(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__host__ __device__ __forceinline__ uint32_t idesc_bf16_major(int M, int N, int a_mn, int b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// one CTA, 128 threads.  K <= 128 rows (multiple of 16), M in {64,128}, N multiple of 16 <= 128.
// raw_out: [128 lanes][N] dump of the TMEM accumulator region.
__global__ void __launch_bounds__(128, 1) umma_probe_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                            int K, int M, int N, float* __restrict__ raw_out) {
    using namespace umma;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t done_bar;
    __shared__ uint32_t tmem_base_s;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int blk = 128 * 128;                                   // one [128 K-rows x 64 cols] block
    unsigned char* sA = smem;                                    // 2 blocks (M up to 128)
    unsigned char* sB = smem + 2 * blk;                          // 2 blocks (N up to 128)
    for (int i = tid; i < 4 * blk / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    __syncthreads();
    for (int i = tid; i < K * M; i += 128) {
        int k = i / M, m = i % M;
        *reinterpret_cast<__nv_bfloat16*>(sA + (m >> 6) * blk + sw128_off(k, m & 63)) = __float2bfloat16_rn(A[i]);
    }
    for (int i = tid; i < K * N; i += 128) {
        int k = i / N, n = i % N;
        *reinterpret_cast<__nv_bfloat16*>(sB + (n >> 6) * blk + sw128_off(k, n & 63)) = __float2bfloat16_rn(B[i]);
    }
    if (tid == 0) { mbar_init(&done_bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&tmem_base_s, 128);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    if (tid == 0) {
        const uint32_t idesc = idesc_bf16_major(M, N, 1, 1);
        for (int k = 0; k < K / 16; ++k) {
            uint64_t da = smem_desc_sw128_mn(smem_u32(sA) + k * 16 * 128, blk);
            uint64_t db = smem_desc_sw128_mn(smem_u32(sB) + k * 16 * 128, blk);
            mma_bf16_ss(tmem_base, da, db, idesc, k > 0 ? 1u : 0u);
        }
        mma_commit(&done_bar);
    }
    mbar_wait(&done_bar, 0);
    tc_fence_after();
    for (int cc = 0; cc < N; cc += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)cc, v);
        for (int j = 0; j < 16; ++j) raw_out[tid * N + cc + j] = v[j];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 128);
}

}  // namespace oevae
