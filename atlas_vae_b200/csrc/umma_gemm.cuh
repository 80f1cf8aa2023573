// Layered tensor-core GEMM: tcgen05.mma (kind::f16, bf16 operands, fp32 accumulate in TMEM) with the operand
// conversion done on load.  Same contract as gemm64_kernel (GemmArgs): fp32 operands with arbitrary strides in
// global memory, fp32 result, fused bias / ReLU / mask / accumulate epilogue, split over the contraction for dW.
//   SPLIT = true  (OEVAE_PREC_BF16X3): x = hi + lo (two bf16), D += Ahi*Bhi + Ahi*Blo + Alo*Bhi  -> ~2^-16 relative
//   SPLIT = false (OEVAE_PREC_BF16)  : operands rounded to bf16 once
// CTA tile: 128 (M, TMEM lanes) x BN (<=128) x 64 (K per stage); 2 smem stages; 256 threads convert+stage the
// operands into the canonical K-major SWIZZLE_128B layout, thread 0 issues the MMAs, warps 0-3 run the epilogue.
#pragma once
#include "umma.cuh"
#include "ref_kernels.cuh"

namespace oevae {

// stage one operand tile [NROWS x 64] (row index = output row/col, 64 contraction steps) as bf16 hi(/lo)
//   element(row, t) = base[row*s_row + t*s_t];  TC: s_t == 1 (contraction contiguous) else s_row == 1
template <bool TC, bool SPLIT>
__device__ __forceinline__ void stage_tile(const float* __restrict__ base, long s_row, long s_t, int nrows_tile,
                                           int row0, int row_end, int ones_row, int tt, int t1,
                                           unsigned char* s_hi, unsigned char* s_lo) {
    using namespace umma;
    const int tid = threadIdx.x;
    if (TC) {
        const int n_vec = nrows_tile * 16;                       // float4 slots along t
        for (int idx = tid; idx < n_vec; idx += 256) {
            int rr = idx >> 4, c4 = (idx & 15) << 2;
            int row = row0 + rr, t = tt + c4;
            float v[4] = {0.f, 0.f, 0.f, 0.f};
            if (row < row_end) {
                if (row == ones_row) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = (t + j < t1) ? 1.f : 0.f;
                } else {
                    const float* p = base + row * s_row + t;
                    if (t + 3 < t1 && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
                        float4 q = __ldg(reinterpret_cast<const float4*>(p));
                        v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) if (t + j < t1) v[j] = __ldg(p + j);
                    }
                }
            }
            __nv_bfloat16 h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) { if (SPLIT) split_bf16(v[j], h[j], l[j]); else h[j] = __float2bfloat16_rn(v[j]); }
            uint32_t off = sw128_off(rr, c4);
            *reinterpret_cast<uint2*>(s_hi + off) = make_uint2(pack2(h[0], h[1]), pack2(h[2], h[3]));
            if (SPLIT) *reinterpret_cast<uint2*>(s_lo + off) = make_uint2(pack2(l[0], l[1]), pack2(l[2], l[3]));
        }
    } else {
        const int rquads = nrows_tile >> 2;                      // 4 consecutive rows per float4
        const int n_task = rquads * 16;                          // x 16 t-quads
        for (int idx = tid; idx < n_task; idx += 256) {
            int rq = idx % rquads, tq = idx / rquads;
            int rr = rq << 2, row = row0 + rr, t = tt + (tq << 2);
            float v[4][4];                                       // [t j][row i]
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
                for (int i = 0; i < 4; ++i) v[j][i] = 0.f;
                if (t + j < t1) {
                    const float* p = base + (long)(t + j) * s_t + row;
                    if (row + 3 < row_end && row + 3 != ones_row && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
                        float4 q = __ldg(reinterpret_cast<const float4*>(p));
                        v[j][0] = q.x; v[j][1] = q.y; v[j][2] = q.z; v[j][3] = q.w;
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (row + i < row_end) v[j][i] = (row + i == ones_row) ? 1.f : __ldg(p + i);
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                __nv_bfloat16 h[4], l[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) { if (SPLIT) split_bf16(v[j][i], h[j], l[j]); else h[j] = __float2bfloat16_rn(v[j][i]); }
                uint32_t off = sw128_off(rr + i, tq << 2);
                *reinterpret_cast<uint2*>(s_hi + off) = make_uint2(pack2(h[0], h[1]), pack2(h[2], h[3]));
                if (SPLIT) *reinterpret_cast<uint2*>(s_lo + off) = make_uint2(pack2(l[0], l[1]), pack2(l[2], l[3]));
            }
        }
    }
}

template <bool A_TC, bool B_TC, bool SPLIT>
__global__ void __launch_bounds__(256, 1) umma_gemm_kernel(GemmArgs g, int BN, int tmem_cols) {
    using namespace umma;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t free_bar[2];
    __shared__ uint64_t done_bar;
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, warp = tid >> 5;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int a_bytes = 128 * 128, b_bytes = BN * 128;
    const int stage_bytes = (a_bytes + b_bytes) * (SPLIT ? 2 : 1);
    auto A_hi = [&](int s) { return smem + s * stage_bytes; };
    auto A_lo = [&](int s) { return smem + s * stage_bytes + a_bytes; };
    auto B_hi = [&](int s) { return smem + s * stage_bytes + (SPLIT ? 2 : 1) * a_bytes; };
    auto B_lo = [&](int s) { return smem + s * stage_bytes + 2 * a_bytes + b_bytes; };

    const int r0 = blockIdx.x * 128, c0 = blockIdx.y * BN;
    int t0 = g.t_base, t1 = g.t_end;
    float* C = g.C;
    if (g.chunk > 0) {
        int grp = (int)blockIdx.z / g.slots_per_group, ck = (int)blockIdx.z - grp * g.slots_per_group;
        t0 = grp * g.group_rows + ck * g.chunk;
        t1 = min(min(t0 + g.chunk, (grp + 1) * g.group_rows), g.t_end);
        C += (long)blockIdx.z * g.slot_stride;
    }
    const int numK = t1 > t0 ? (t1 - t0 + 63) >> 6 : 0;

    if (tid == 0) { mbar_init(&free_bar[0], 1); mbar_init(&free_bar[1], 1); mbar_init(&done_bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&tmem_base_s, (uint32_t)tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = idesc_bf16(128, BN);
    const int onesA = g.ones_row ? g.R - 1 : -1;

    for (int kb = 0; kb < numK; ++kb) {
        const int s = kb & 1, tt = t0 + (kb << 6);
        if (kb >= 2) mbar_wait(&free_bar[s], (uint32_t)(((kb >> 1) - 1) & 1));
        stage_tile<A_TC, SPLIT>(g.A, g.sAr, g.sAt, 128, r0, g.R, onesA, tt, t1, A_hi(s), A_lo(s));
        stage_tile<B_TC, SPLIT>(g.B, g.sBc, g.sBt, BN, c0, g.Cn, -1, tt, t1, B_hi(s), B_lo(s));
        fence_async_smem();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t a_hi = smem_u32(A_hi(s)), b_hi = smem_u32(B_hi(s));
            const int ksteps = min(4, (t1 - tt + 15) >> 4);
            for (int k = 0; k < ksteps; ++k) {
                uint64_t da = smem_desc_sw128(a_hi + k * 32), db = smem_desc_sw128(b_hi + k * 32);
                mma_bf16_ss(tmem_base, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                if (SPLIT) {
                    uint64_t dal = smem_desc_sw128(smem_u32(A_lo(s)) + k * 32), dbl = smem_desc_sw128(smem_u32(B_lo(s)) + k * 32);
                    mma_bf16_ss(tmem_base, da, dbl, idesc, 1u);
                    mma_bf16_ss(tmem_base, dal, db, idesc, 1u);
                }
            }
            mma_commit(&free_bar[s]);
            if (kb == numK - 1) mma_commit(&done_bar);
        }
    }
    if (numK > 0) { mbar_wait(&done_bar, 0); tc_fence_after(); }

    // ---- epilogue: TMEM lane = output row; warps 0-3 cover the 128 lanes
    if (warp < 4) {
        const int r = r0 + tid;
        for (int cc = 0; cc < BN; cc += 16) {
            float v[16];
            if (numK > 0) tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)cc, v);
            else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
            }
            if (r < g.R) {
                float* crow = C + (long)r * g.ldc;
#pragma unroll
                for (int j0 = 0; j0 < 16; j0 += 4) {
                    int c = c0 + cc + j0;
                    if (c >= g.Cn) break;
                    float o[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float x = v[j0 + j];
                        if (c + j < g.Cn) {
                            if (g.bias) x += __ldg(g.bias + c + j);
                            if (g.accumulate) x += crow[c + j];
                            if (g.mask) x = (__ldg(g.mask + (long)r * g.ldmask + c + j) > 0.f) ? x : 0.f;
                            if (g.relu) x = fmaxf(x, 0.f);
                        }
                        o[j] = x;
                    }
                    if (c + 3 < g.Cn && ((reinterpret_cast<uintptr_t>(crow + c) & 15) == 0)) {
                        *reinterpret_cast<float4*>(crow + c) = make_float4(o[0], o[1], o[2], o[3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) if (c + j < g.Cn) crow[c + j] = o[j];
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, (uint32_t)tmem_cols);
}

}  // namespace oevae
