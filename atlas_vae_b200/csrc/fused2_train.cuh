// Feature-major fused OE-VAE step for nets whose layers fit one CTA (plan: fused2_plan.h).
//   forward + losses  models.py:30-36,49-52,62-69,72-85,123-132      gradient of the batch SUM  models.py:183
// Every Dense contraction (forward, dX chain, dW^T) is a tcgen05.mma (kind::f16, bf16 hi+lo split operands = 3 MMAs,
// fp32 accumulators in TMEM) with the layer's FEATURES on the 128 TMEM lanes and the tile's 64 jets on the columns.
// Activations never leave shared memory; weights stream from the pre-swizzled bf16 image (adam_img_kernel) through a
// 4-slot cp.async.bulk ring; one CTA = a strided set of 64-jet tiles; its dW/db go to a private slice of the
// partial-gradient buffer (summed in fixed order by adam_img_kernel).
//
// TMEM columns: [0,64) mu (kept until the latent backward), [64,128) log_var, [128,192) chain accumulator,
// [192,256) / [256,320) output-layer chunk accumulators, [320,448) dW^T accumulator; dW_0^T reuses [0,448).
#pragma once
#include "umma.cuh"
#include "umma_probe.cuh"
#include "fused2_plan.h"
#include "common.cuh"

namespace oevae {

namespace f2 {
using namespace umma;

constexpr uint32_t T_MU = 0, T_LV = 64, T_CH = 128, T_Y0 = 192, T_DW = 320;

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// one MMA operand in shared memory: hi part address, distance to the lo part, bytes per 16-deep k-step,
// MN-major flag (+ leading byte offset between 64-element MN blocks)
struct Opnd { uint32_t hi, lo, kstep, lbo; int mn; };
__device__ __forceinline__ Opnd opnd_k(uint32_t hi, uint32_t lo) { return Opnd{hi, lo, 32u, 0u, 0}; }              // K-major, K <= 64
__device__ __forceinline__ Opnd opnd_mn(uint32_t hi, uint32_t lo, uint32_t lbo) { return Opnd{hi, lo, 2048u, lbo, 1}; }
__device__ __forceinline__ uint64_t desc(const Opnd& o, int ks, bool lo) {
    const uint32_t a = o.hi + (lo ? o.lo : 0u) + (uint32_t)ks * o.kstep;
    return o.mn ? smem_desc_sw128_mn(a, o.lbo) : smem_desc_sw128(a);
}
// D[128 x N] (+)= A * B over nk k-steps, bf16x3: Ahi*Bhi + Ahi*Blo + Alo*Bhi
__device__ __forceinline__ void issue(uint32_t d, const Opnd& A, const Opnd& B, int nk, int N, bool acc) {
    const uint32_t idesc = idesc_bf16_major(128, N, A.mn, B.mn);
    for (int ks = 0; ks < nk; ++ks) {
        const uint64_t ah = desc(A, ks, false), al = desc(A, ks, true), bh = desc(B, ks, false), bl = desc(B, ks, true);
        mma_bf16_ss(d, ah, bh, idesc, acc ? 1u : 0u);
        mma_bf16_ss(d, ah, bl, idesc, 1u);
        mma_bf16_ss(d, al, bh, idesc, 1u);
        acc = true;
    }
}

// 16 consecutive jets (col0 % 16 == 0) of feature row `row` of an activation tile, as bf16 hi/lo
__device__ __forceinline__ void store16(unsigned char* t_hi, unsigned char* t_lo, int row, int col0, const float* v) {
    uint32_t h[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        __nv_bfloat16 h0, l0, h1, l1;
        split_bf16(v[2 * j], h0, l0); split_bf16(v[2 * j + 1], h1, l1);
        h[j] = pack2(h0, h1); l[j] = pack2(l0, l1);
    }
    const uint32_t o0 = sw128_off(row, col0), o1 = sw128_off(row, col0 + 8);
    *reinterpret_cast<uint4*>(t_hi + o0) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(t_hi + o1) = make_uint4(h[4], h[5], h[6], h[7]);
    *reinterpret_cast<uint4*>(t_lo + o0) = make_uint4(l[0], l[1], l[2], l[3]);
    *reinterpret_cast<uint4*>(t_lo + o1) = make_uint4(l[4], l[5], l[6], l[7]);
}
// bit j set <=> hi part of (row, col0 + j) is > 0
__device__ __forceinline__ uint32_t positive16(const unsigned char* t_hi, int row, int col0) {
    const uint4 q0 = *reinterpret_cast<const uint4*>(t_hi + sw128_off(row, col0));
    const uint4 q1 = *reinterpret_cast<const uint4*>(t_hi + sw128_off(row, col0 + 8));
    const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t lo16 = w[j] & 0xffffu, hi16 = w[j] >> 16;
        if (lo16 != 0u && !(lo16 & 0x8000u)) m |= 1u << (2 * j);
        if (hi16 != 0u && !(hi16 & 0x8000u)) m |= 1u << (2 * j + 1);
    }
    return m;
}
// v[i] of every lane summed over the 32 lanes; lane i returns the total of index i
__device__ __forceinline__ float transpose_reduce32(float* v, int lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const bool up = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const float send = up ? v[i] : v[i + s], keep = up ? v[i + s] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    return v[0];
}
}  // namespace f2

struct F2Args {
    const float* x; const float* w; const float* eps;      // eps nullable (then Philox)
    const float* theta;                                     // flat fp32 parameters (biases are read from here)
    const unsigned char* wimg; float* parts;                // parts: [gridDim.x][Ppad]
    float* per_jet; double* sums;                           // per_jet nullable [4,B]
    long B; int n_tiles; int train; int mse;
    float beta;
    unsigned long long seed, row_offset; unsigned int step;
    long long* dbg;
};

__global__ void __launch_bounds__(256, 1) fused2_kernel(const __grid_constant__ F2Plan P, F2Args a) {
    using namespace umma;
    using namespace f2;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t mma_bar, y_bar, xbar[kF2XSlots], wbar[kF2WSlots];
    __shared__ uint32_t tmem_base_s;
    unsigned char* sm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = warp & 3, h = warp >> 2;
    const int frow = 32 * q + lane;                          // TMEM lane = feature row owned by this thread
    const int jc0 = 32 * h;                                  // first of the 32 jet columns owned by this thread
    const int n = P.n, D = P.D, L = P.L, so = 2 * n + 1;
    const int nops = a.train ? P.n_op_train : P.n_op_eval;
    const int n_fwd_ops = P.op_first[so] + P.nX;

    float* s_w = reinterpret_cast<float*>(sm + P.scratch_off);        // [64] jet weights (0 for rows past the batch)
    float* s_rec = s_w + 64;                                           // [4][64] per lane-quarter reconstruction sums
    float* s_kld = s_rec + 256;                                        // [64]
    float* s_db = s_kld + 64;                                          // [4][2][128] bias-gradient halves
    float* s_eps = s_db + 1024;                                        // [L][64]
    auto tile_hi = [&](int k) { return sm + P.tile_off[k]; };
    auto tile_lo = [&](int k) { return sm + P.tile_off[k] + P.tile_rows[k] * 128; };
    auto xslot = [&](int s) { return sm + P.s_off + s * 16384; };      // hi 8 KB, lo 8 KB
    unsigned char* gy_hi = sm + P.s_off; unsigned char* gy_lo = gy_hi + 16384;
    auto wslot = [&](long g) { return sm + P.wring_off + (int)(g % kF2WSlots) * P.wslot_bytes; };

    // ---- one-time setup
    if (tid == 0) {
        mbar_init(&mma_bar, 1); mbar_init(&y_bar, 1);
        for (int i = 0; i < kF2XSlots; ++i) mbar_init(&xbar[i], 1);
        for (int i = 0; i < kF2WSlots; ++i) mbar_init(&wbar[i], 1);
        mbar_fence_init();
    }
    for (int i = tid * 16; i < P.zero_bytes; i += 256 * 16) *reinterpret_cast<uint4*>(sm + i) = make_uint4(0, 0, 0, 0);
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    const uint32_t t_lane = tm + ((uint32_t)(32 * q) << 16);

    // ---- weight ring (thread 0 only): ops are consumed strictly in order; op g lives in slot g % 4
    const int my_tiles = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const long w_total = (long)my_tiles * nops;
    long w_issued = 0, w_done = 0, gbase = 0;
    int w_next = 0;
    auto refill = [&]() {
        while (w_issued < w_total && w_issued < w_done + kF2WSlots) {
            const int k = (int)(w_issued % nops), s = (int)(w_issued % kF2WSlots);
            const uint32_t bytes = (!a.train && k >= n_fwd_ops) ? 16u : P.op_bytes[k];
            mbar_expect_tx(&wbar[s], bytes);
            bulk_g2s(wslot(w_issued), a.wimg + P.op_off[k], bytes, &wbar[s]);
            ++w_issued;
        }
    };
    auto wget = [&](int k) -> uint32_t {                      // wait until op k of this tile has landed; returns its slot
        for (; w_next <= k; ++w_next) { const long g = gbase + w_next; mbar_wait(&wbar[g % kF2WSlots], (uint32_t)((g / kF2WSlots) & 1)); }
        return smem_u32(wslot(gbase + k));
    };
    auto wdone = [&](int k) { w_done = gbase + k + 1; refill(); };   // every MMA that reads ops <= k has completed
    if (tid == 0) refill();

    uint32_t ph_m = 0, ph_y = 0, ph_x = 0;
    auto wait_mma = [&]() { mbar_wait(&mma_bar, ph_m); ph_m ^= 1u; tc_fence_after(); };
    auto wait_y = [&]() { mbar_wait(&y_bar, ph_y); ph_y ^= 1u; tc_fence_after(); };
    auto wait_x = [&](int s) { mbar_wait(&xbar[s], (ph_x >> s) & 1u); ph_x ^= 1u << s; tc_fence_after(); };
    auto publish = [&]() { fence_async_smem(); tc_fence_before(); __syncthreads(); };   // smem writes + TMEM reads done

    int n_dbg = 0;
    auto stamp = [&]() { if (a.dbg && blockIdx.x == 0 && tid == 0 && n_dbg < 255) a.dbg[1 + n_dbg++] = clock64(); };
    stamp();

    int tile_iter = 0;
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++tile_iter) {
        const long row0 = (long)tile * kF2Jets;
        const int nvalid = (int)min((long)kF2Jets, a.B - row0);
        float* part = a.parts + (long)blockIdx.x * P.Ppad;
        w_next = 0;
        if (tid < 64) s_w[tid] = tid < nvalid ? a.w[row0 + tid] : 0.f;
        for (int pf = tile_iter ? 1 : 0; pf < 2; ++pf) {   // pull the next tile's rows of X (and the first tile's own) towards L2
            const long nrow0 = row0 + (long)pf * gridDim.x * kF2Jets;
            const long nbytes = min((long)kF2Jets, a.B - nrow0) * D * 4;
            const char* nx = reinterpret_cast<const char*>(a.x + nrow0 * D);
            for (long o = (long)tid * 128; o < nbytes; o += 256 * 128) prefetch_l2(nx + o);
        }

        // [64 jets x features 64c..64c+63] of X -> ring slot xs (bf16 hi/lo, K-major rows = jets)
        auto stage_x = [&](int c, int xs) {
            unsigned char* hi = xslot(xs); unsigned char* lo = hi + 8192;
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int idx = tid + 256 * it, r = idx >> 4, c4 = (idx & 15) << 2, col = 64 * c + c4;
                float v[4] = {0.f, 0.f, 0.f, 0.f};
                if (r < nvalid) {
                    const float* p = a.x + (row0 + r) * D + col;
                    if (col + 3 < D && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
                        const float4 t = __ldg(reinterpret_cast<const float4*>(p));
                        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) v[j] = (col + j < D) ? __ldg(p + j) : 0.f;
                    }
                }
                __nv_bfloat16 hh[4], ll[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) split_bf16(v[j], hh[j], ll[j]);
                const uint32_t off = sw128_off(r, c4);
                *reinterpret_cast<uint2*>(hi + off) = make_uint2(pack2(hh[0], hh[1]), pack2(hh[2], hh[3]));
                *reinterpret_cast<uint2*>(lo + off) = make_uint2(pack2(ll[0], ll[1]), pack2(ll[2], ll[3]));
            }
        };
        // chain accumulator -> bias + ReLU -> tile s (pad rows zero)
        auto epi_act = [&](int s) {
            const F2Stage& st = P.st[s];
            if (32 * q >= P.tile_rows[s]) return;
            const float bias = frow < st.out ? __ldg(a.theta + st.seg_b[0] + frow) : 0.f;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                float v[16];
                tmem_ld16(t_lane + T_CH + jc0 + 16 * g, v);
                if (frow < P.tile_rows[s]) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float t = v[j] + bias;
                        v[j] = frow < st.out ? (st.relu ? fmaxf(t, 0.f) : t) : 0.f;
                    }
                    store16(tile_hi(s), tile_lo(s), frow, jc0 + 16 * g, v);
                }
            }
        };
        // chain accumulator (dX of stage s) -> ReLU mask of H_{s-1} -> G_{s-1} over tile s-1, bias-gradient halves
        auto epi_grad = [&](int s) {
            const F2Stage& st = P.st[s];
            const int rows = P.tile_rows[s - 1];
            if (32 * q >= rows) return;
            float dbs = 0.f;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                float v[16];
                tmem_ld16(t_lane + T_CH + jc0 + 16 * g, v);
                if (frow < rows) {
                    const uint32_t m = positive16(tile_hi(s - 1), frow, jc0 + 16 * g);
#pragma unroll
                    for (int j = 0; j < 16; ++j) { v[j] = (((m >> j) & 1u) && frow < st.in) ? v[j] : 0.f; dbs += v[j]; }
                    store16(tile_hi(s - 1), tile_lo(s - 1), frow, jc0 + 16 * g, v);
                }
            }
            if (frow < rows) s_db[(((s - 1) & 1) * 2 + h) * 128 + frow] = dbs;
        };
        // dW^T accumulator of stage s (lane = output feature, column = input feature) -> partial-gradient slice
        auto epi_dw = [&](int s, int buf) {
            const F2Stage& st = P.st[s];
            const int top = st.seg_lane0[st.nseg - 1] + st.seg_n[st.nseg - 1];
            if (32 * q >= top) return;
            int sg = -1;
#pragma unroll
            for (int k = 0; k < 2; ++k) if (k < st.nseg && frow >= st.seg_lane0[k] && frow < st.seg_lane0[k] + st.seg_n[k]) sg = k;
            const int o = sg >= 0 ? frow - st.seg_lane0[sg] : 0, ld = sg >= 0 ? st.seg_n[sg] : 0;
            float* dw = part + (sg >= 0 ? st.seg_w[sg] : 0) + o;
            const int ng = st.kin >> 4, g0 = h ? (ng + 1) >> 1 : 0, g1 = h ? ng : (ng + 1) >> 1;
            for (int g = g0; g < g1; ++g) {
                float v[16];
                tmem_ld16(t_lane + T_DW + 16 * g, v);
                if (sg >= 0) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int i = 16 * g + j;
                        if (i < st.in) { float* d = dw + (long)i * ld; *d = tile_iter ? *d + v[j] : v[j]; }
                    }
                }
            }
            if (sg >= 0 && h == 0) {
                float* d = part + st.seg_b[sg] + o;
                const float t = s_db[(buf * 2) * 128 + frow] + s_db[(buf * 2 + 1) * 128 + frow];
                *d = tile_iter ? *d + t : t;
            }
        };

        // =========================================================== forward, stage 0: X streamed in 64-feature chunks
        {
            const F2Stage& st = P.st[0];
            for (int c = 0; c < P.nX; ++c) {
                const int xs = c % kF2XSlots;
                if (c >= kF2XSlots) { wait_x(xs); if (tid == 0) wdone(P.op_first[0] + c - kF2XSlots); }
                stage_x(c, xs);
                publish();
                if (tid == 0) {
                    const uint32_t wb = wget(P.op_first[0] + c);
                    tc_fence_after();
                    const int nk = min(4, (D - 64 * c + 15) >> 4);
                    issue(tm + T_CH, opnd_k(wb, st.rows * 128), opnd_k(smem_u32(xslot(xs)), 8192), nk, kF2Jets, c > 0);
                    mma_commit(&xbar[xs]);
                }
            }
            for (int c = max(0, P.nX - kF2XSlots); c < P.nX; ++c) wait_x(c % kF2XSlots);
            if (tid == 0) wdone(P.op_first[0] + P.nX - 1);
            stamp();
            epi_act(0);
            publish();
            stamp();
        }
        // =========================================================== forward chain: stages 1 .. 2n
        for (int s = 1; s <= 2 * n; ++s) {
            const F2Stage& st = P.st[s];
            if (tid == 0) {
                const uint32_t wb = wget(P.op_first[s]);
                tc_fence_after();
                const uint32_t part_b = st.rows * 128;
                const Opnd B = opnd_mn(smem_u32(tile_hi(s - 1)), P.tile_rows[s - 1] * 128, 8192);
                const int nk = st.kin >> 4;
                if (st.kind == 1) issue(tm + T_CH, opnd_mn(wb, part_b, part_b), B, nk, kF2Jets, false);
                else if (st.kind == 2) issue(tm + T_CH, opnd_k(wb, part_b), B, nk, kF2Jets, false);
                else {                                                   // heads: mean block, log_var block
                    issue(tm + T_MU, opnd_mn(wb, 2 * part_b, part_b), B, nk, kF2Jets, false);
                    issue(tm + T_LV, opnd_mn(wb + part_b, 2 * part_b, part_b), B, nk, kF2Jets, false);
                }
                mma_commit(&mma_bar);
            }
            wait_mma();
            if (tid == 0) wdone(P.op_first[s]);
            if (s != n) {
                epi_act(s);
            } else if (q == 0) {
                // ---- heads: Sampling (models.py:9-15) + KLD (models.py:81-85); z -> rows 0..15 of the latent tile
                const F2Stage& hs = P.st[n];
                const float b_mu = lane < L ? __ldg(a.theta + hs.seg_b[0] + lane) : 0.f;
                const float b_lv = lane < L ? __ldg(a.theta + hs.seg_b[1] + lane) : 0.f;
                float kl[32];
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    float mu[16], lv[16], ep[16], z[16];
                    tmem_ld16(t_lane + T_MU + jc0 + 16 * g, mu);
                    tmem_ld16(t_lane + T_LV + jc0 + 16 * g, lv);
                    if (a.eps) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int col = jc0 + 16 * g + j;
                            ep[j] = (lane < L && col < nvalid) ? __ldg(a.eps + (row0 + col) * L + lane) : 0.f;
                        }
                    } else {
                        // Philox counter (row, quad, step) yields latents 4*quad..4*quad+3 of one jet: lane (quad, t) draws
                        // jet cb+t and the four lanes of a quad exchange components
#pragma unroll
                        for (int cb = 0; cb < 16; cb += 4) {
                            float nrm[4];
                            philox_normal4(a.seed, a.row_offset + (unsigned long long)(row0 + jc0 + 16 * g + cb + (lane & 3)),
                                           (unsigned)(lane >> 2), a.step, nrm);
#pragma unroll
                            for (int r = 0; r < 4; ++r) {
                                const int sel = (lane & 3) ^ r;
                                const float pub = sel == 0 ? nrm[0] : sel == 1 ? nrm[1] : sel == 2 ? nrm[2] : nrm[3];
                                const float got = __shfl_sync(0xffffffffu, pub, (lane & ~3) | sel);
                                // got = component (lane & 3) of jet cb + sel
#pragma unroll
                                for (int t = 0; t < 4; ++t) if (t == sel) ep[cb + t] = got;
                            }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int col = jc0 + 16 * g + j;
                        float zz = 0.f, kk = 0.f;
                        if (lane < L) {
                            const float m = mu[j] + b_mu, v = lv[j] + b_lv;
                            zz = m + clipv(expf(v * 0.5f)) * ep[j];
                            kk = 1.f + v - clipv(expf(v)) - m * m;
                            s_eps[lane * 64 + col] = ep[j];
                        }
                        z[j] = zz; kl[16 * g + j] = kk;
                    }
                    if (lane < 16) store16(tile_hi(n), tile_lo(n), lane, jc0 + 16 * g, z);
                }
                const float ks = transpose_reduce32(kl, lane);
                s_kld[jc0 + lane] = -(ks / (float)L) / 2.f;
            }
            publish();
            stamp();
        }
        // =========================================================== output layer in 128-feature chunks
        float rec_acc[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) rec_acc[j] = 0.f;
        {
            const F2Stage& st = P.st[so];
            const uint32_t part_b = st.rows * 128;
            const int ob = P.op_first[so];
            const float inv_d = 1.f / (float)D;
            auto issue_y = [&](int c) {                                 // thread 0: y chunk c = W_out[:, chunk]^T d_n
                wget(ob + min(2 * c + 1, P.nX - 1));
                const uint32_t wb = smem_u32(wslot(gbase + ob + 2 * c));
                tc_fence_after();
                issue(tm + T_Y0 + 64 * (c & 1), opnd_mn(wb, part_b, P.wslot_bytes),
                      opnd_mn(smem_u32(tile_hi(2 * n)), P.tile_rows[2 * n] * 128, 8192), st.kin >> 4, kF2Jets, false);
                mma_commit(&y_bar);
            };
            if (tid == 0) { wget(ob - 1); wdone(ob - 1); issue_y(0); }   // ob-1: stage 2n's op or the alignment dummy
            for (int c = 0; c < P.nC; ++c) {
                const int last_op = ob + min(2 * c + 1, P.nX - 1);
                wait_y();
                if (tid == 0 && c + 1 < P.nC) issue_y(c + 1);
                stamp();
                // reconstruction loss (models.py:66,72-78) and its gradient through clip_values
                {
                    const int f = 128 * c + frow;
                    const bool valid = f < D;
                    const float bias = valid ? __ldg(a.theta + st.seg_b[0] + f) : 0.f;
                    float dbs = 0.f;
#pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        float y[16], xv[16];
                        tmem_ld16(t_lane + T_Y0 + 64 * (c & 1) + jc0 + 16 * g, y);
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int col = jc0 + 16 * g + j;
                            xv[j] = (valid && col < nvalid) ? __ldg(a.x + (row0 + col) * D + f) : 0.f;
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            float gy = 0.f;
                            if (valid) {
                                const float yr = y[j] + bias, d = clipv(yr) - xv[j], wj = s_w[jc0 + 16 * g + j];
                                rec_acc[16 * g + j] += a.mse ? d * d : fabsf(d);
                                const float gg = a.mse ? wj * 2.f * d * inv_d : wj * (float)((d > 0.f) - (d < 0.f)) * inv_d;
                                gy = clip_pass(yr) ? gg : 0.f;
                            }
                            dbs += gy; y[j] = gy;
                        }
                        if (a.train) store16(gy_hi, gy_lo, frow, jc0 + 16 * g, y);
                    }
                    if (a.train) s_db[((2 + (c & 1)) * 2 + h) * 128 + frow] = dbs;
                }
                stamp();
                if (!a.train) {
                    tc_fence_before();
                    __syncthreads();                                    // accumulator reads done before chunk c+2 overwrites it
                    if (tid == 0) wdone(last_op);
                    continue;
                }
                publish();
                if (tid == 0) {
                    tc_fence_after();
                    // dW_out^T[chunk] = g_y (A: K-major, K = jets) x d_n (B: K-major, N = in features)
                    issue(tm + T_DW, opnd_k(smem_u32(gy_hi), 16384), opnd_k(smem_u32(tile_hi(2 * n)), P.tile_rows[2 * n] * 128),
                          4, st.kin, false);
                    // g_dn += W_out[:, chunk] x g_y[chunk]   (A: K-major rows = in features; B: MN-major, K = chunk features)
                    for (int b = 0; b < 2 && 2 * c + b < P.nX; ++b) {
                        const uint32_t wb = wget(ob + 2 * c + b);
                        const int nk = min(4, (D - 64 * (2 * c + b) + 15) >> 4);
                        issue(tm + T_CH, opnd_k(wb, part_b), opnd_mn(smem_u32(gy_hi) + b * 8192, 16384, 8192), nk, kF2Jets, c > 0 || b > 0);
                    }
                    mma_commit(&mma_bar);
                }
                wait_mma();
                if (tid == 0) wdone(last_op);
                stamp();
                // dW_out^T chunk -> part (lanes = output features: coalesced rows of the Keras [in, D] kernel)
                {
                    const int f = 128 * c + frow;
                    const bool valid = f < D;
                    float* dw = part + st.seg_w[0] + f;
                    const int ng = st.kin >> 4, g0 = h ? (ng + 1) >> 1 : 0, g1 = h ? ng : (ng + 1) >> 1;
                    for (int g = g0; g < g1; ++g) {
                        float v[16];
                        tmem_ld16(t_lane + T_DW + 16 * g, v);
                        if (valid) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int i = 16 * g + j;
                                if (i < st.in) { float* d = dw + (long)i * D; *d = tile_iter ? *d + v[j] : v[j]; }
                            }
                        }
                    }
                    if (valid && h == 0) {
                        float* d = part + st.seg_b[0] + f;
                        const float t = s_db[((2 + (c & 1)) * 2) * 128 + frow] + s_db[((2 + (c & 1)) * 2 + 1) * 128 + frow];
                        *d = tile_iter ? *d + t : t;
                    }
                }
                tc_fence_before();
                stamp();
            }
            const float rs = transpose_reduce32(rec_acc, lane);
            s_rec[q * 64 + jc0 + lane] = rs;
            if (a.train) { __syncthreads(); epi_grad(so); publish(); }
            stamp();
        }
        // =========================================================== backward chain: stages 2n .. 1
        if (a.train) {
            for (int s = 2 * n; s >= 1; --s) {
                const F2Stage& st = P.st[s];
                if (tid == 0) {
                    const uint32_t wb = wget(P.op_bwd[s]);
                    tc_fence_after();
                    const uint32_t part_b = st.rows * 128;
                    const uint32_t g_hi = smem_u32(tile_hi(s)), g_lo = P.tile_rows[s] * 128;
                    // dW_s^T = G_s (A: K-major, K = jets) x H_{s-1} (B: K-major, N = in features)
                    issue(tm + T_DW, opnd_k(g_hi, g_lo), opnd_k(smem_u32(tile_hi(s - 1)), P.tile_rows[s - 1] * 128), 4, st.kin, false);
                    // G_{s-1}^T = W_s x G_s^T
                    const Opnd G = opnd_mn(g_hi, g_lo, 8192);
                    if (st.kind == 1) issue(tm + T_CH, opnd_k(wb, part_b), G, st.kout >> 4, kF2Jets, false);
                    else if (st.kind == 2) issue(tm + T_CH, opnd_mn(wb, part_b, part_b), G, st.kout >> 4, kF2Jets, false);
                    else {
                        issue(tm + T_CH, opnd_k(wb, 2 * part_b), G, 1, kF2Jets, false);                                      // W_mu g_mu
                        issue(tm + T_CH, opnd_k(wb + part_b, 2 * part_b), opnd_mn(g_hi + 2048, g_lo, 8192), 1, kF2Jets, true);   // + W_lv g_lv
                    }
                    mma_commit(&mma_bar);
                }
                wait_mma();
                if (tid == 0) wdone(P.op_bwd[s]);
                epi_dw(s, s & 1);
                if (s != n + 1) {
                    epi_grad(s);
                } else if (q == 0) {
                    // ---- latent backward: g_z -> g_mu (rows 0..15), g_logvar (rows 16..31) of the latent tile
                    const F2Stage& hs = P.st[n];
                    const float b_mu = lane < L ? __ldg(a.theta + hs.seg_b[0] + lane) : 0.f;
                    const float b_lv = lane < L ? __ldg(a.theta + hs.seg_b[1] + lane) : 0.f;
                    const float inv_l = 1.f / (float)L;
                    float dbm = 0.f, dbl = 0.f;
#pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        float gz[16], mu[16], lv[16];
                        tmem_ld16(t_lane + T_CH + jc0 + 16 * g, gz);
                        tmem_ld16(t_lane + T_MU + jc0 + 16 * g, mu);
                        tmem_ld16(t_lane + T_LV + jc0 + 16 * g, lv);
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int col = jc0 + 16 * g + j;
                            float gm = 0.f, gl = 0.f;
                            if (lane < L) {
                                const float m = mu[j] + b_mu, v = lv[j] + b_lv, gk = a.beta * s_w[col];
                                const float e_raw = expf(v), s_raw = expf(v * 0.5f);
                                gm = gz[j] + gk * m * inv_l;
                                gl = gk * (-(1.f - (clip_pass(e_raw) ? e_raw : 0.f)) * 0.5f * inv_l)
                                   + gz[j] * s_eps[lane * 64 + col] * (clip_pass(s_raw) ? s_raw * 0.5f : 0.f);
                            }
                            mu[j] = gm; lv[j] = gl; dbm += gm; dbl += gl;
                        }
                        if (lane < 16) {
                            store16(tile_hi(n), tile_lo(n), lane, jc0 + 16 * g, mu);
                            store16(tile_hi(n), tile_lo(n), 16 + lane, jc0 + 16 * g, lv);
                        }
                    }
                    if (lane < 16) { s_db[((n & 1) * 2 + h) * 128 + lane] = dbm; s_db[((n & 1) * 2 + h) * 128 + 16 + lane] = dbl; }
                }
                publish();
                stamp();
            }
            // ======================================================= stage 0: dW_0^T = G_0 x X, X re-staged from L2
            {
                const F2Stage& st = P.st[0];
                for (int c0 = 0; c0 < P.nX; c0 += 7) {
                    const int c1 = min(P.nX, c0 + 7);
                    for (int c = c0; c < c1; ++c) {
                        const int xs = (c - c0) % kF2XSlots;
                        if (c - c0 >= kF2XSlots) wait_x(xs);
                        stage_x(c, xs);
                        publish();
                        if (tid == 0) {
                            tc_fence_after();
                            issue(tm + 64 * (c - c0), opnd_k(smem_u32(tile_hi(0)), P.tile_rows[0] * 128),
                                  opnd_mn(smem_u32(xslot(xs)), 8192, 8192), 4, 64, false);
                            mma_commit(&xbar[xs]);
                        }
                    }
                    for (int c = max(c0, c1 - kF2XSlots); c < c1; ++c) wait_x((c - c0) % kF2XSlots);
                    stamp();
                    if (32 * q < st.out) {
                        const bool valid = frow < st.out;
                        float* dw = part + st.seg_w[0] + frow;
                        const int ng = (c1 - c0) * 4, g0 = h ? (ng + 1) >> 1 : 0, g1 = h ? ng : (ng + 1) >> 1;
                        for (int g = g0; g < g1; ++g) {
                            float v[16];
                            tmem_ld16(t_lane + 16 * g, v);
                            if (valid) {
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    const int i = 64 * c0 + 16 * g + j;
                                    if (i < D) { float* d = dw + (long)i * st.out; *d = tile_iter ? *d + v[j] : v[j]; }
                                }
                            }
                        }
                        if (valid && h == 0 && c0 == 0) {
                            float* d = part + st.seg_b[0] + frow;
                            const float t = s_db[frow] + s_db[128 + frow];
                            *d = tile_iter ? *d + t : t;
                        }
                    }
                    publish();
                }
            }
        }
        stamp();
        // =========================================================== per-jet losses (models.py:123-132) and batch sums
        if (tid == 0) { wget(nops - 1); wdone(nops - 1); gbase += nops; }
        __syncthreads();
        if (tid < 64) {
            const int j = tid;
            float lr_ = 0.f, lk = 0.f, tot = 0.f;
            if (j < nvalid) {
                const float rec = (s_rec[j] + s_rec[64 + j] + s_rec[128 + j] + s_rec[192 + j]) / (float)D, wj = s_w[j];
                lr_ = rec * wj; lk = s_kld[j] * wj * a.beta; tot = lr_ + lk + 0.f;
                if (a.per_jet) {
                    a.per_jet[row0 + j] = lr_; a.per_jet[a.B + row0 + j] = lk;
                    a.per_jet[2 * a.B + row0 + j] = 0.f; a.per_jet[3 * a.B + row0 + j] = tot;
                }
            }
            if (a.sums) {
                const double v0 = warp_sum((double)lr_), v1 = warp_sum((double)lk), v3 = warp_sum((double)tot);
                if (lane == 0) { atomicAdd(a.sums, v0); atomicAdd(a.sums + 1, v1); atomicAdd(a.sums + 3, v3); }
            }
        }
        __syncthreads();
    }
    stamp();
    if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[0] = n_dbg;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}

}  // namespace oevae
