// Probe kernel (tests only): validates the MN-major SWIZZLE_128B descriptors and the M=64 / M=128 TMEM accumulator
// layouts that the fused training kernel relies on.  C[M,N] = A^T B with A [K, M] (M contiguous), B [K, N]
// (N contiguous), both staged as [K rows x 64-column blocks] (the activation-tile layout) and consumed MN-major.
#pragma once
#include "umma.cuh"

namespace oevae {

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

namespace oevae {
// Latency probe (tests/tuning only): cycles from the issue of `nm` back-to-back tcgen05.mma (same accumulator, K-major
// operands) to the wake-up of the mbarrier wait that follows the commit.  out[i] for nm = 1, 3, 6, 12, 18, 36.
__global__ void __launch_bounds__(128, 1) umma_latency_kernel(int M, int N, long long* out) {
    using namespace umma;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 2 * 128 * 128 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&tmem_base_s, 256);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    uint32_t ph = 0;
    const int nms[6] = {1, 3, 6, 12, 18, 36};
    if (tid == 0) {
        const uint32_t idesc = idesc_bf16(M, N);
        for (int t = 0; t < 6; ++t) {
            long long best = 1ll << 60;
            for (int rep = 0; rep < 5; ++rep) {
                long long c0 = clock64();
                for (int k = 0; k < nms[t]; ++k)
                    mma_bf16_ss(tm, smem_desc_sw128(smem_u32(smem) + (k & 3) * 32), smem_desc_sw128(smem_u32(smem) + 128 * 128 + (k & 3) * 32), idesc, k > 0);
                mma_commit(&bar);
                mbar_wait(&bar, ph); ph ^= 1u;
                long long c1 = clock64();
                if (c1 - c0 < best) best = c1 - c0;
            }
            out[t] = best;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 256);
}
}  // namespace oevae

namespace oevae {
// Issue-rate probe: `nw` warps (1..4) each issue `nm` MMAs (one elected lane per warp, own accumulator, own barrier);
// mode 0: descriptors rebuilt per MMA, mode 1: descriptors hoisted (only the 32-bit low word advances).
// out[0] = cycles until every warp has seen its commit.
__global__ void __launch_bounds__(128, 1) umma_issue_kernel(int M, int N, int nw, int nm, int mode, long long* out) {
    using namespace umma;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t tmem_base_s;
    __shared__ long long t_end[4];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 2 * 128 * 128 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    if (tid == 0) { for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    const uint32_t idesc = idesc_bf16(M, N);
    long long best = 1ll << 60;
    for (int rep = 0; rep < 5; ++rep) {
        __syncthreads();
        long long c0 = clock64();
        if (warp < nw && lane == 0) {
            const uint32_t d = tm + warp * 128;
            if (mode == 0) {
                for (int k = 0; k < nm; ++k)
                    mma_bf16_ss(d, smem_desc_sw128(smem_u32(smem) + (k & 3) * 32), smem_desc_sw128(smem_u32(smem) + 128 * 128 + (k & 3) * 32), idesc, k > 0);
            } else if (mode == 1) {
                const uint64_t da = smem_desc_sw128(smem_u32(smem)), db = smem_desc_sw128(smem_u32(smem) + 128 * 128);
#pragma unroll 4
                for (int k = 0; k < nm; ++k) mma_bf16_ss(d, da + (uint64_t)((k & 3) * 2), db + (uint64_t)((k & 3) * 2), idesc, k > 0);
            } else {
                // modes 2..4: MN-major operands (64-element blocks of [64 k x 64 mn], 8 KB apart): 2 = B, 3 = A and B, 4 = A
                const int a_mn = mode == 3 || mode == 4, b_mn = mode == 2 || mode == 3;
                const uint32_t id2 = idesc_bf16_major(M, N, a_mn, b_mn);
                const uint32_t sa = smem_u32(smem), sb = smem_u32(smem) + 128 * 128;
                const uint64_t da = a_mn ? smem_desc_sw128_mn(sa, 8192) : smem_desc_sw128(sa), db = b_mn ? smem_desc_sw128_mn(sb, 8192) : smem_desc_sw128(sb);
                const uint64_t sta = a_mn ? (2048 >> 4) : 2, stb = b_mn ? (2048 >> 4) : 2;
#pragma unroll 4
                for (int k = 0; k < nm; ++k) mma_bf16_ss(d, da + (uint64_t)(k & 3) * sta, db + (uint64_t)(k & 3) * stb, id2, k > 0);
            }
            mma_commit(&bar[warp]);
            mbar_wait(&bar[warp], (uint32_t)(rep & 1));
            t_end[warp] = clock64();
        }
        __syncthreads();
        long long c1 = 0;
        for (int i = 0; i < nw; ++i) c1 = max(c1, t_end[i]);
        if (c1 - c0 < best) best = c1 - c0;
    }
    if (tid == 0) out[0] = best;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}
}  // namespace oevae

namespace oevae {
// Bulk-copy probe: cycles for `depth` cp.async.bulk copies of `bytes` each (global, L2-resident after the first rep)
// issued back to back by one thread until the last mbarrier completes.  out[0] best-of-5, out[1] first (cold) rep.
__global__ void __launch_bounds__(128, 1) bulk_copy_probe_kernel(const unsigned char* src, int bytes, int depth, long cta_stride, long long* out) {
    using namespace umma;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t bar[8];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x;
    if (tid == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bar[i], 1); mbar_fence_init(); }
    __syncthreads();
    if (tid == 0) {
        long long best = 1ll << 60, first = 0;
        for (int rep = 0; rep < 5; ++rep) {
            long long c0 = clock64();
            for (int i = 0; i < depth; ++i) {
                asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(&bar[i])), "r"((uint32_t)bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(smem + (size_t)i * bytes)), "l"(src + (size_t)blockIdx.x * cta_stride + (size_t)i * bytes), "r"((uint32_t)bytes), "r"(smem_u32(&bar[i])) : "memory");
            }
            for (int i = 0; i < depth; ++i) mbar_wait(&bar[i], (uint32_t)(rep & 1));
            long long c1 = clock64();
            if (rep == 0) first = c1 - c0;
            if (c1 - c0 < best) best = c1 - c0;
        }
        if (blockIdx.x == 0) { out[0] = best; out[1] = first; }
    }
}
}  // namespace oevae
