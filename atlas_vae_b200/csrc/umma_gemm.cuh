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

constexpr int kGT = 512;                                    // threads of the GEMM CTA (16 warps: staging needs the issue slots and the latency hiding)
constexpr int kGS = 2048 / kGT;                              // float4 slots per thread and operand tile (128 rows x 16 quads)
constexpr int kGK = kGS / 4;                                 // !TC tasks per thread (4 slots each)

// One operand tile [nrows_tile x 64] (row index = output row/col, 64 contraction steps) goes global -> registers -> shared
// memory in two phases, so that all of a thread's loads are in flight together and the NEXT k-block's loads can be issued
// before the current one is converted: a 256-thread CTA that loads, converts and stores one float4 at a time spends
// 8 DRAM latencies per tile, which is what bounded this kernel (15 k cycles per k-block against 0.8 k of MMA time).
//   element(row, t) = base[row*s_row + t*s_t];  TC: s_t == 1 (contraction contiguous) else s_row == 1
// TC  : slot i of a thread = float4 along t of row (tid + kGT i) >> 4
// !TC : slots 4i..4i+3 = float4 along rows (4 consecutive rows) at t, t+1, t+2, t+3 of task tid + kGT i
template <bool TC>
__device__ __forceinline__ void load_tile(const float* __restrict__ base, long s_row, long s_t, int nrows_tile,
                                          int row0, int row_end, int ones_row, int tt, int t1, float4 (&r)[kGS]) {
    const int tid = threadIdx.x;
    if (TC) {
        const int n_vec = nrows_tile * 16;                       // float4 slots along t
#pragma unroll
        for (int i = 0; i < kGS; ++i) {
            const int idx = tid + kGT * i;
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
        for (int i = 0; i < kGK; ++i) {
            const int idx = tid + kGT * i;
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
// Slot numbering is load_tile's.  TC: slot i = row (tid >> 4) + (kGT / 16) i, t-quad tid & 15.  !TC: task i = tid + 256 i -> row-quad rq,
// t-quad tq; slot 4 i + j = rows 4 rq .. 4 rq + 3 at t = 4 tq + j.
struct StageGeo {
    int n;                   // TC: slots of this thread; !TC: tasks of this thread
    int g0, gstep;           // TC: element offset of slot 0 in the tile, step per slot
    uint32_t s0;             // TC: byte offset of slot 0 in the smem tile (slot i: + 2048 i -- the swizzle term only sees row & 7)
    int gt[kGK]; uint32_t st[kGK], cx[kGK];   // !TC, per task: element offset of (t = 4 tq, row = 4 rq); smem offset of the k = 4 tq row + column bytes; swizzle chunk
};
template <bool TC>
__device__ __forceinline__ StageGeo stage_geo(long s_row, long s_t, int nrows_tile) {
    using namespace umma;
    const int tid = threadIdx.x;
    StageGeo G{};
    if (TC) {
        G.n = max(0, (nrows_tile * 16 - tid + kGT - 1) / kGT);     // slot i is valid iff tid + kGT i < 16 nrows
        G.g0 = (tid >> 4) * (int)s_row + ((tid & 15) << 2); G.gstep = (kGT / 16) * (int)s_row;
        G.s0 = sw128_off(tid >> 4, (tid & 15) << 2);
    } else {
        const int rquads = nrows_tile >> 2, n_task = rquads * 16;
        G.n = 0;
#pragma unroll
        for (int i = 0; i < kGK; ++i) {
            const int idx = tid + kGT * i;
            if (idx < n_task) {
                G.n = i + 1;
                const int rq = idx % rquads, tq = idx / rquads, rr = rq << 2, mn = rr & 63;
                G.gt[i] = (tq << 2) * (int)s_t + rr;
                G.st[i] = (uint32_t)((rr >> 6) * 8192 + (tq << 2) * 128 + ((mn & 7) << 1));
                G.cx[i] = (uint32_t)((mn >> 3) ^ ((tq & 1) << 2));
            }
        }
    }
    return G;
}
// whole tile inside the matrix, no ones row, every float4 aligned
template <bool TC>
__device__ __forceinline__ bool stage_fast(const float* base, long s_row, long s_t, int nrows_tile, int row0, int row_end, int ones_row, int t0) {
    const bool inside = row0 + nrows_tile <= row_end && (ones_row < row0 || ones_row >= row0 + nrows_tile);
    if (TC) return inside && (s_row & 3) == 0 && ((reinterpret_cast<uintptr_t>(base + (long)row0 * s_row + t0) & 15) == 0);
    return inside && (s_t & 3) == 0 && ((reinterpret_cast<uintptr_t>(base + (long)t0 * s_t + row0) & 15) == 0);
}
// interior tile: every slot is one aligned float4; `tile` = first element of the tile at this k-block
template <bool TC>
__device__ __forceinline__ void load_fast(const float* __restrict__ tile, long s_t, const StageGeo& G, float4 (&r)[kGS]) {
    if (TC) {
#pragma unroll
        for (int i = 0; i < kGS; ++i) r[i] = i < G.n ? __ldg(reinterpret_cast<const float4*>(tile + G.g0 + i * G.gstep)) : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
#pragma unroll
        for (int i = 0; i < kGK; ++i) {
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
__device__ __forceinline__ void store_tile(const float4 (&r)[kGS], const StageGeo& G, unsigned char* s_hi, unsigned char* s_lo) {
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
        for (int i = 0; i < kGS; ++i) if (i < G.n) put(G.s0 + (uint32_t)(kGT / 16 * 128) * i, r[i]);
    } else {
#pragma unroll
        for (int i = 0; i < kGK; ++i) {
            if (i < G.n) {
#pragma unroll
                for (int j = 0; j < 4; ++j) put(G.st[i] + 128u * j + ((G.cx[i] ^ (uint32_t)j) << 4), r[4 * i + j]);
            }
        }
    }
}

// Persistent: one CTA per SM walks the (slot, row block, column block) tiles with the column block fastest, so the CTAs that share
// an A tile run together (A comes from HBM once, from L2 for the rest; a layer's weights stay in L2 anyway).  TMEM, the barriers
// and the staging geometry are set up once; the first k-block of the next tile is loaded into registers before the epilogue of
// the current one, so that neither CTA launch nor the first DRAM latency of a tile is exposed.
// Epilogue: TMEM -> shared memory (row pitch BN + 4 floats) -> global in whole rows of the tile (coalesced 16-byte accesses for
// the result, the mask and the accumulate input; a lane-per-row epilogue touched 32 lines per instruction).
template <bool A_TC, bool B_TC, bool SPLIT>
__global__ void __launch_bounds__(kGT, 1) umma_gemm_kernel(GemmArgs g, int BN, int tmem_cols, int n_slots) {
    using namespace umma;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t free_bar[2];
    __shared__ uint64_t done_bar;
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int a_bytes = 128 * 128, b_bytes = B_TC ? BN * 128 : ((BN + 63) >> 6) * 8192;
    const int stage_bytes = (a_bytes + b_bytes) * (SPLIT ? 2 : 1);
    auto A_hi = [&](int s) { return smem + s * stage_bytes; };
    auto A_lo = [&](int s) { return smem + s * stage_bytes + a_bytes; };
    auto B_hi = [&](int s) { return smem + s * stage_bytes + (SPLIT ? 2 : 1) * a_bytes; };
    auto B_lo = [&](int s) { return smem + s * stage_bytes + 2 * a_bytes + b_bytes; };
    float* s_out = reinterpret_cast<float*>(smem);          // epilogue staging [128][BN + 4] (the operand stages are idle then)
    const int opitch = BN + 4;

    if (tid == 0) { mbar_init(&free_bar[0], 1); mbar_init(&free_bar[1], 1); mbar_init(&done_bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&tmem_base_s, (uint32_t)tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = idesc_bf16_major(128, BN, A_TC ? 0 : 1, B_TC ? 0 : 1);
    const int onesA = g.ones_row ? g.R - 1 : -1;
    const int n_cb = (g.Cn + BN - 1) / BN, n_rb = (g.R + 127) >> 7;
    const int n_tiles = n_cb * n_rb * n_slots;

    struct Tile { int r0, c0, t0, t1, numK, slot; bool fa, fb; };
    auto tile_at = [&](int tile) {
        Tile T;
        T.slot = tile / (n_cb * n_rb);
        const int rc = tile - T.slot * (n_cb * n_rb);
        T.r0 = (rc / n_cb) * 128; T.c0 = (rc % n_cb) * BN;
        T.t0 = g.t_base; T.t1 = g.t_end;
        if (g.chunk > 0) {
            const int grp = T.slot / g.slots_per_group, ck = T.slot - grp * g.slots_per_group;
            T.t0 = grp * g.group_rows + ck * g.chunk;
            T.t1 = min(min(T.t0 + g.chunk, (grp + 1) * g.group_rows), g.t_end);
        }
        T.numK = T.t1 > T.t0 ? (T.t1 - T.t0 + 63) >> 6 : 0;
        T.fa = stage_fast<A_TC>(g.A, g.sAr, g.sAt, 128, T.r0, g.R, onesA, T.t0);
        T.fb = stage_fast<B_TC>(g.B, g.sBc, g.sBt, BN, T.c0, g.Cn, -1, T.t0);
        return T;
    };
    const StageGeo GA = stage_geo<A_TC>(g.sAr, g.sAt, 128);
    const StageGeo GB = stage_geo<B_TC>(g.sBc, g.sBt, BN);
    // operands of k-block tt of tile T -> registers: aligned float4 loads off the precomputed offsets when the tile is interior
    // and the k-block is whole, the bounds-checked loader otherwise (the register layout is the same)
    auto load_a = [&](const Tile& T, int tt, float4 (&r)[kGS]) {
        if (T.fa && tt + 64 <= T.t1) load_fast<A_TC>(A_TC ? g.A + (long)T.r0 * g.sAr + tt : g.A + (long)tt * g.sAt + T.r0, g.sAt, GA, r);
        else load_tile<A_TC>(g.A, g.sAr, g.sAt, 128, T.r0, g.R, onesA, tt, T.t1, r);
    };
    auto load_b = [&](const Tile& T, int tt, float4 (&r)[kGS]) {
        if (T.fb && tt + 64 <= T.t1) load_fast<B_TC>(B_TC ? g.B + (long)T.c0 * g.sBc + tt : g.B + (long)tt * g.sBt + T.c0, g.sBt, GB, r);
        else load_tile<B_TC>(g.B, g.sBc, g.sBt, BN, T.c0, g.Cn, -1, tt, T.t1, r);
    };

    uint32_t gk = 0, nd = 0;                                 // k-blocks staged so far (stage = gk & 1), done_bar phases consumed
    float4 ra[kGS], rb[kGS], na[kGS], nb[kGS];
    int tile = blockIdx.x;
    Tile T = tile < n_tiles ? tile_at(tile) : Tile{};
    if (tile < n_tiles && T.numK > 0) { load_a(T, T.t0, ra); load_b(T, T.t0, rb); }
    for (; tile < n_tiles; tile += gridDim.x) {
        const int ntile = tile + gridDim.x;
        const Tile N = ntile < n_tiles ? tile_at(ntile) : Tile{};
        // one k-block: `ca, cb` hold its operands; the next k-block's (or the next tile's first) leave for `xa, xb` before
        // these are converted.  The caller alternates the two register sets, so nothing is copied.
        auto kblock = [&](int kb, float4 (&ca)[kGS], float4 (&cb)[kGS], float4 (&xa)[kGS], float4 (&xb)[kGS]) {
            const int s = gk & 1, tt = T.t0 + (kb << 6);
            const bool more = kb + 1 < T.numK, next_tile = !more && ntile < n_tiles && N.numK > 0;
            if (more) { load_a(T, tt + 64, xa); load_b(T, tt + 64, xb); }
            else if (next_tile) { load_a(N, N.t0, xa); load_b(N, N.t0, xb); }
            if (gk >= 2) mbar_wait(&free_bar[s], (uint32_t)(((gk >> 1) - 1) & 1));
            store_tile<A_TC, SPLIT>(ca, GA, A_hi(s), A_lo(s));
            store_tile<B_TC, SPLIT>(cb, GB, B_hi(s), B_lo(s));
            fence_async_smem();
            __syncthreads();
            if (warp == 0) {                                       // warp-uniform issue: elect.sync picks the lane (see umma.cuh mma_e)
                tc_fence_after();
                const uint32_t a_hi = smem_u32(A_hi(s)), b_hi = smem_u32(B_hi(s)), a_lo = smem_u32(A_lo(s)), b_lo = smem_u32(B_lo(s));
                const int ksteps = min(4, (T.t1 - tt + 15) >> 4);
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
                if (kb == T.numK - 1) commit_e(&done_bar);
            }
            ++gk;
        };
        for (int kb = 0; kb < T.numK; kb += 2) {
            kblock(kb, ra, rb, na, nb);
            if (kb + 1 < T.numK) kblock(kb + 1, na, nb, ra, rb);
        }
        if (T.numK & 1) {                                        // the next tile's first k-block sits in the other register set
#pragma unroll
            for (int i = 0; i < kGS; ++i) { ra[i] = na[i]; rb[i] = nb[i]; }
        }
        if (T.numK == 0 && ntile < n_tiles && N.numK > 0) { load_a(N, N.t0, ra); load_b(N, N.t0, rb); }
        if (T.numK > 0) { mbar_wait(&done_bar, nd & 1); ++nd; tc_fence_after(); }

        // ---- epilogue, phase 1: TMEM lane = output row; warp w reads lane quarter w & 3, 16-column groups go round the warp quads; rows go to shared memory (all MMAs that read the operand stages have completed)
        for (int cc = 16 * (warp >> 2); cc < BN; cc += 16 * (kGT / 128)) {
            float v[16];
            if (T.numK > 0) tmem_ld16(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)cc, v);
            else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.f;
            }
            float* srow = s_out + (tid & 127) * opitch + cc;
#pragma unroll
            for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(srow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        tc_fence_before();
        __syncthreads();
        // ---- phase 2: a warp writes whole rows (lane = 4 consecutive columns); bias / accumulate / mask / ReLU on the way
        {
            float* C = g.C + (g.chunk > 0 ? (long)T.slot * g.slot_stride : 0);
            const int nrow = min(128, g.R - T.r0);
            for (int cq = lane * 4; cq < BN; cq += 128) {
                const int c = T.c0 + cq;
                if (c >= g.Cn) break;
                const bool full = c + 3 < g.Cn;
                float b4[4] = {0.f, 0.f, 0.f, 0.f};
                if (g.bias) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) if (c + j < g.Cn) b4[j] = __ldg(g.bias + c + j);
                }
                for (int rr = warp; rr < nrow; rr += kGT / 32) {
                    const long r = T.r0 + rr;
                    const float4 q = *reinterpret_cast<const float4*>(s_out + rr * opitch + cq);
                    float o[4] = {q.x + b4[0], q.y + b4[1], q.z + b4[2], q.w + b4[3]};
                    float* cp = g.C ? C + r * g.ldc + c : nullptr;
                    const bool vec = full && ((reinterpret_cast<uintptr_t>(cp) & 15) == 0);
                    if (g.accumulate) {
                        if (vec) { const float4 p = *reinterpret_cast<const float4*>(cp); o[0] += p.x; o[1] += p.y; o[2] += p.z; o[3] += p.w; }
                        else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) if (c + j < g.Cn) o[j] += cp[j];
                        }
                    }
                    if (g.mask) {
                        const float* mp = g.mask + r * g.ldmask + c;
                        if (full && ((reinterpret_cast<uintptr_t>(mp) & 15) == 0)) {
                            const float4 m = __ldg(reinterpret_cast<const float4*>(mp));
                            o[0] = m.x > 0.f ? o[0] : 0.f; o[1] = m.y > 0.f ? o[1] : 0.f; o[2] = m.z > 0.f ? o[2] : 0.f; o[3] = m.w > 0.f ? o[3] : 0.f;
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) if (c + j < g.Cn) o[j] = __ldg(mp + j) > 0.f ? o[j] : 0.f;
                        }
                    }
                    if (g.relu) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) o[j] = fmaxf(o[j], 0.f);
                    }
                    if (vec) *reinterpret_cast<float4*>(cp) = make_float4(o[0], o[1], o[2], o[3]);
                    else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) if (c + j < g.Cn) cp[j] = o[j];
                    }
                }
            }
        }
        __syncthreads();                                         // s_out is the operand stages of the next tile
        T = N;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, (uint32_t)tmem_cols);
}

}  // namespace oevae
