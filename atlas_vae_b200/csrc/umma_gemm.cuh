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

// One operand tile [nrows_tile x 64] (row index = output row/col, 64 contraction steps) goes global -> registers -> shared
// memory in two phases, so that all of a thread's loads are in flight together and the NEXT k-block's loads can be issued
// before the current one is converted: a 256-thread CTA that loads, converts and stores one float4 at a time spends
// 8 DRAM latencies per tile, which is what bounded this kernel (15 k cycles per k-block against 0.8 k of MMA time).
//   element(row, t) = base[row*s_row + t*s_t];  TC: s_t == 1 (contraction contiguous) else s_row == 1
// TC  : slot i of a thread = float4 along t of row (tid + 256 i) >> 4
// !TC : slots 4i..4i+3 = float4 along rows (4 consecutive rows) at t, t+1, t+2, t+3 of task tid + 256 i
template <bool TC>
__device__ __forceinline__ void load_tile(const float* __restrict__ base, long s_row, long s_t, int nrows_tile,
                                          int row0, int row_end, int ones_row, int tt, int t1, float4 (&r)[8]) {
    const int tid = threadIdx.x;
    if (TC) {
        const int n_vec = nrows_tile * 16;                       // float4 slots along t
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = tid + 256 * i;
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < n_vec) {
                const int rr = idx >> 4, c4 = (idx & 15) << 2;
                const int row = row0 + rr, t = tt + c4;
                if (row < row_end) {
                    if (row == ones_row) {
                        q.x = (t < t1) ? 1.f : 0.f; q.y = (t + 1 < t1) ? 1.f : 0.f; q.z = (t + 2 < t1) ? 1.f : 0.f; q.w = (t + 3 < t1) ? 1.f : 0.f;
                    } else {
                        const float* p = base + row * s_row + t;
                        if (t + 3 < t1 && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) q = __ldg(reinterpret_cast<const float4*>(p));
                        else {
                            if (t < t1) q.x = __ldg(p);
                            if (t + 1 < t1) q.y = __ldg(p + 1);
                            if (t + 2 < t1) q.z = __ldg(p + 2);
                            if (t + 3 < t1) q.w = __ldg(p + 3);
                        }
                    }
                }
            }
            r[i] = q;
        }
    } else {
        const int rquads = nrows_tile >> 2;                      // 4 consecutive rows per float4
        const int n_task = rquads * 16;                          // x 16 t-quads
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int idx = tid + 256 * i;
            const bool on = idx < n_task;
            const int rq = on ? idx % rquads : 0, tq = on ? idx / rquads : 0;
            const int row = row0 + (rq << 2), t = tt + (tq << 2);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                if (on && t + j < t1) {
                    const float* p = base + (long)(t + j) * s_t + row;
                    if (row + 3 < row_end && row + 3 != ones_row && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) q = __ldg(reinterpret_cast<const float4*>(p));
                    else {
                        if (row < row_end) q.x = (row == ones_row) ? 1.f : __ldg(p);
                        if (row + 1 < row_end) q.y = (row + 1 == ones_row) ? 1.f : __ldg(p + 1);
                        if (row + 2 < row_end) q.z = (row + 2 == ones_row) ? 1.f : __ldg(p + 2);
                        if (row + 3 < row_end) q.w = (row + 3 == ones_row) ? 1.f : __ldg(p + 3);
                    }
                }
                r[4 * i + j] = q;
            }
        }
    }
}

// Per-thread staging geometry of one operand; nothing in it depends on the k-block, so the k loop only adds the tile base.
// Slot numbering is load_tile's.  TC: slot i = row (tid >> 4) + 16 i, t-quad tid & 15.  !TC: task i = tid + 256 i -> row-quad rq,
// t-quad tq; slot 4 i + j = rows 4 rq .. 4 rq + 3 at t = 4 tq + j.
struct StageGeo {
    int n;                   // TC: slots of this thread; !TC: tasks of this thread
    int g0, gstep;           // TC: element offset of slot 0 in the tile, step per slot
    uint32_t s0;             // TC: byte offset of slot 0 in the smem tile (slot i: + 2048 i -- the swizzle term only sees row & 7)
    int gt[2]; uint32_t st[2], cx[2];   // !TC, per task: element offset of (t = 4 tq, row = 4 rq); smem offset of the k = 4 tq row + column bytes; swizzle chunk
    bool fast;               // whole tile inside the matrix, no ones row, every float4 aligned
};
template <bool TC>
__device__ __forceinline__ StageGeo stage_geo(const float* base, long s_row, long s_t, int nrows_tile, int row0, int row_end, int ones_row, int t0) {
    using namespace umma;
    const int tid = threadIdx.x;
    StageGeo G{};
    const bool inside = row0 + nrows_tile <= row_end && (ones_row < row0 || ones_row >= row0 + nrows_tile);
    if (TC) {
        G.n = nrows_tile >> 4;
        G.g0 = (tid >> 4) * (int)s_row + ((tid & 15) << 2); G.gstep = 16 * (int)s_row;
        G.s0 = sw128_off(tid >> 4, (tid & 15) << 2);
        G.fast = inside && (s_row & 3) == 0 && ((reinterpret_cast<uintptr_t>(base + (long)row0 * s_row + t0) & 15) == 0);
    } else {
        const int rquads = nrows_tile >> 2, n_task = rquads * 16;
        G.n = 0;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int idx = tid + 256 * i;
            if (idx < n_task) {
                G.n = i + 1;
                const int rq = idx % rquads, tq = idx / rquads, rr = rq << 2, mn = rr & 63;
                G.gt[i] = (tq << 2) * (int)s_t + rr;
                G.st[i] = (uint32_t)((rr >> 6) * 8192 + (tq << 2) * 128 + ((mn & 7) << 1));
                G.cx[i] = (uint32_t)((mn >> 3) ^ ((tq & 1) << 2));
            }
        }
        G.fast = inside && (s_t & 3) == 0 && ((reinterpret_cast<uintptr_t>(base + (long)t0 * s_t + row0) & 15) == 0);
    }
    return G;
}
// interior tile: every slot is one aligned float4; `tile` = first element of the tile at this k-block
template <bool TC>
__device__ __forceinline__ void load_fast(const float* __restrict__ tile, long s_t, const StageGeo& G, float4 (&r)[8]) {
    if (TC) {
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = i < G.n ? __ldg(reinterpret_cast<const float4*>(tile + G.g0 + i * G.gstep)) : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                r[4 * i + j] = i < G.n ? __ldg(reinterpret_cast<const float4*>(tile + G.gt[i] + j * (int)s_t)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
}
// registers -> bf16 hi(/lo) in shared memory, SWIZZLE_128B.  TC operands (contraction contiguous in global memory) land
// K-major: [rows x 64 t].  Row-contiguous operands land MN-major: 64-row blocks of [64 t x 64 rows] (8 KB each), so a float4 of
// four consecutive rows stays four consecutive bf16 -- no transpose through shared memory (the K-major placement of these
// operands made every store a 16-way bank conflict) -- and the MMA reads them through an MN-major descriptor.
template <bool TC, bool SPLIT>
__device__ __forceinline__ void store_tile(const float4 (&r)[8], const StageGeo& G, unsigned char* s_hi, unsigned char* s_lo) {
    using namespace umma;
    auto put = [&](uint32_t off, const float4& q) {
        if (SPLIT) {
            uint32_t h0, l0, h1, l1;
            split_pair(q.x, q.y, h0, l0); split_pair(q.z, q.w, h1, l1);
            *reinterpret_cast<uint2*>(s_hi + off) = make_uint2(h0, h1);
            *reinterpret_cast<uint2*>(s_lo + off) = make_uint2(l0, l1);
        } else {
            *reinterpret_cast<uint2*>(s_hi + off) = make_uint2(round_pair(q.x, q.y), round_pair(q.z, q.w));
        }
    };
    if (TC) {
#pragma unroll
        for (int i = 0; i < 8; ++i) if (i < G.n) put(G.s0 + 2048u * i, r[i]);
    } else {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            if (i < G.n) {
#pragma unroll
                for (int j = 0; j < 4; ++j) put(G.st[i] + 128u * j + ((G.cx[i] ^ (uint32_t)j) << 4), r[4 * i + j]);
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

    const int tid = threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int a_bytes = 128 * 128, b_bytes = B_TC ? BN * 128 : ((BN + 63) >> 6) * 8192;
    const int stage_bytes = (a_bytes + b_bytes) * (SPLIT ? 2 : 1);
    auto A_hi = [&](int s) { return smem + s * stage_bytes; };
    auto A_lo = [&](int s) { return smem + s * stage_bytes + a_bytes; };
    auto B_hi = [&](int s) { return smem + s * stage_bytes + (SPLIT ? 2 : 1) * a_bytes; };
    auto B_lo = [&](int s) { return smem + s * stage_bytes + 2 * a_bytes + b_bytes; };

    // blockIdx.x = row block x column blocks with the column block fastest: the CTAs that share an A tile run together, so A
    // comes from HBM once and from L2 for the rest (the weights of a layer stay in L2 anyway)
    const int n_cb = (g.Cn + BN - 1) / BN;
    const int r0 = (int)(blockIdx.x / n_cb) * 128, c0 = (int)(blockIdx.x % n_cb) * BN;
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
    const uint32_t idesc = idesc_bf16_major(128, BN, A_TC ? 0 : 1, B_TC ? 0 : 1);
    const int onesA = g.ones_row ? g.R - 1 : -1;

    const StageGeo GA = stage_geo<A_TC>(g.A, g.sAr, g.sAt, 128, r0, g.R, onesA, t0);
    const StageGeo GB = stage_geo<B_TC>(g.B, g.sBc, g.sBt, BN, c0, g.Cn, -1, t0);
    // operands of k-block tt -> registers: aligned float4 loads off precomputed offsets when the tile is interior and the
    // k-block is whole, the bounds-checked loader otherwise (the register layout is the same)
    auto load_a = [&](int tt, float4 (&r)[8]) {
        if (GA.fast && tt + 64 <= t1) load_fast<A_TC>(A_TC ? g.A + (long)r0 * g.sAr + tt : g.A + (long)tt * g.sAt + r0, g.sAt, GA, r);
        else load_tile<A_TC>(g.A, g.sAr, g.sAt, 128, r0, g.R, onesA, tt, t1, r);
    };
    auto load_b = [&](int tt, float4 (&r)[8]) {
        if (GB.fast && tt + 64 <= t1) load_fast<B_TC>(B_TC ? g.B + (long)c0 * g.sBc + tt : g.B + (long)tt * g.sBt + c0, g.sBt, GB, r);
        else load_tile<B_TC>(g.B, g.sBc, g.sBt, BN, c0, g.Cn, -1, tt, t1, r);
    };
    float4 ra[8], rb[8];
    if (numK > 0) { load_a(t0, ra); load_b(t0, rb); }
    for (int kb = 0; kb < numK; ++kb) {
        const int s = kb & 1, tt = t0 + (kb << 6);
        float4 na[8], nb[8];                                     // the next k-block's operands leave for the SM before this one is converted
        if (kb + 1 < numK) { load_a(tt + 64, na); load_b(tt + 64, nb); }
        if (kb >= 2) mbar_wait(&free_bar[s], (uint32_t)(((kb >> 1) - 1) & 1));
        store_tile<A_TC, SPLIT>(ra, GA, A_hi(s), A_lo(s));
        store_tile<B_TC, SPLIT>(rb, GB, B_hi(s), B_lo(s));
        if (kb + 1 < numK) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { ra[i] = na[i]; rb[i] = nb[i]; }
        }
        fence_async_smem();
        __syncthreads();
        if (warp == 0) {                                       // warp-uniform issue: elect.sync picks the lane (see umma.cuh mma_e)
            tc_fence_after();
            const uint32_t a_hi = smem_u32(A_hi(s)), b_hi = smem_u32(B_hi(s)), a_lo = smem_u32(A_lo(s)), b_lo = smem_u32(B_lo(s));
            const int ksteps = min(4, (t1 - tt + 15) >> 4);
            for (int k = 0; k < ksteps; ++k) {
                // one 16-deep k-step: 32 B along a K-major row, 16 rows (2 KB) of an MN-major block
                auto dsc_a = [&](uint32_t base) { return A_TC ? smem_desc_sw128(base + k * 32) : smem_desc_sw128_mn(base + k * 2048, 8192); };
                auto dsc_b = [&](uint32_t base) { return B_TC ? smem_desc_sw128(base + k * 32) : smem_desc_sw128_mn(base + k * 2048, 8192); };
                const uint64_t da = dsc_a(a_hi), db = dsc_b(b_hi);
                mma_e(tmem_base, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                if (SPLIT) {
                    mma_e(tmem_base, da, dsc_b(b_lo), idesc, 1u);
                    mma_e(tmem_base, dsc_a(a_lo), db, idesc, 1u);
                }
            }
            commit_e(&free_bar[s]);
            if (kb == numK - 1) commit_e(&done_bar);
        }
    }
    if (numK > 0) { mbar_wait(&done_bar, 0); tc_fence_after(); }

    // ---- epilogue: TMEM lane = output row; warp w reads lane quarter w & 3, warps 0-3 take the even 16-column groups and
    // warps 4-7 the odd ones
    {
        const int r = r0 + (tid & 127);
        for (int cc = 16 * (warp >> 2); cc < BN; cc += 32) {
            float v[16];
            if (numK > 0) tmem_ld16(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)cc, v);
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
