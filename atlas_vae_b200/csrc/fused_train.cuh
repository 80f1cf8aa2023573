// Fused per-tile OE-VAE step (models.py:108-132 forward + losses, models.py:183 gradient of the batch SUM) for nets
// whose layers fit one CTA: every Dense contraction of the forward pass, of the dX chain and of the dW/db products
// runs on tcgen05.mma (kind::f16, bf16 hi+lo split operands = 3 MMAs, fp32 accumulators in TMEM); activations never
// leave shared memory; weights stream from pre-swizzled bf16 images (written by adam_img_kernel) with
// cp.async.bulk + mbarrier one operation ahead.  One CTA = one tile of Rt jets; its dW/db go to a private slice of
// the partial-gradient buffer (deterministic; summed by adam_img_kernel).
//
// Data layout (see fused_plan.h): activation tile = [Rt x 64-col blocks] bf16 SWIZZLE_128B, hi part then lo part.
// The same bytes serve as K-major A (forward / dX: M = jets) and as MN-major A or B (dW: K = jets).
// TMEM (512 columns): [0,128) forward / dX accumulator (M=64: jet i -> lane (i%16)+32*(i/16)),
//                     [128,192) output-layer chunk, [256,384) dW accumulator (lane = input feature).
#pragma once
#include "umma.cuh"
#include "umma_probe.cuh"
#include "fused_plan.h"
#include "common.cuh"

namespace oevae {

namespace fz {
using namespace umma;

constexpr uint32_t ACC_F = 0, ACC_Y = 128, ACC_DW = 256;

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// D (+)= A*B^T with hi/lo split operands: Ahi*Bhi + Ahi*Blo + Alo*Bhi
__device__ __forceinline__ void mma3(uint32_t d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t idesc, bool acc) {
    mma_bf16_ss(d, ah, bh, idesc, acc ? 1u : 0u);
    mma_bf16_ss(d, ah, bl, idesc, 1u);
    mma_bf16_ss(d, al, bh, idesc, 1u);
}

// K-major x K-major group: A blocks a_stride apart (lo part at +a_lo), B K-block tiles b_stride apart (lo at +b_lo);
// K = (n_kb-1)*64 + 16*last_ks
__device__ __forceinline__ void issue_kk(uint32_t d, uint32_t a, uint32_t a_lo, uint32_t a_stride, uint32_t b, uint32_t b_lo,
                                         uint32_t b_stride, int n_kb, int last_ks, uint32_t idesc, bool acc_first) {
    bool acc = acc_first;
    for (int kb = 0; kb < n_kb; ++kb) {
        const int ns = (kb == n_kb - 1) ? last_ks : 4;
        for (int ks = 0; ks < ns; ++ks) {
            uint32_t ao = a + kb * a_stride + ks * 32, bo = b + kb * b_stride + ks * 32;
            mma3(d, smem_desc_sw128(ao), smem_desc_sw128(ao + a_lo), smem_desc_sw128(bo), smem_desc_sw128(bo + b_lo), idesc, acc);
            acc = true;
        }
    }
}
// MN-major x MN-major group (dW): rows of both operands are the jets (K), ksteps = Rt/16
__device__ __forceinline__ void issue_mn(uint32_t d, uint32_t a, uint32_t a_lo, uint32_t b, uint32_t b_lo, uint32_t lbo,
                                         int ksteps, uint32_t idesc) {
    for (int ks = 0; ks < ksteps; ++ks) {
        uint32_t ao = a + ks * 2048, bo = b + ks * 2048;
        mma3(d, smem_desc_sw128_mn(ao, lbo), smem_desc_sw128_mn(ao + a_lo, lbo), smem_desc_sw128_mn(bo, lbo),
             smem_desc_sw128_mn(bo + b_lo, lbo), idesc, ks > 0);
    }
}

// store 16 consecutive columns (col0 % 16 == 0) of one row of a tile as bf16 hi/lo
__device__ __forceinline__ void store16(unsigned char* t_hi, unsigned char* t_lo, int blk, int row, int col0, const float* v) {
    uint32_t h[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        __nv_bfloat16 h0, l0, h1, l1;
        split_bf16(v[2 * j], h0, l0); split_bf16(v[2 * j + 1], h1, l1);
        h[j] = pack2(h0, h1); l[j] = pack2(l0, l1);
    }
    const int b = col0 >> 6, c = col0 & 63;
    uint32_t o0 = b * blk + sw128_off(row, c), o1 = b * blk + sw128_off(row, c + 8);
    *reinterpret_cast<uint4*>(t_hi + o0) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(t_hi + o1) = make_uint4(h[4], h[5], h[6], h[7]);
    *reinterpret_cast<uint4*>(t_lo + o0) = make_uint4(l[0], l[1], l[2], l[3]);
    *reinterpret_cast<uint4*>(t_lo + o1) = make_uint4(l[4], l[5], l[6], l[7]);
}
// sign pattern (> 0) of 16 consecutive columns of the hi part
__device__ __forceinline__ uint32_t positive16(const unsigned char* t_hi, int blk, int row, int col0) {
    const int b = col0 >> 6, c = col0 & 63;
    uint4 q0 = *reinterpret_cast<const uint4*>(t_hi + b * blk + sw128_off(row, c));
    uint4 q1 = *reinterpret_cast<const uint4*>(t_hi + b * blk + sw128_off(row, c + 8));
    uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t lo16 = w[j] & 0xffffu, hi16 = w[j] >> 16;
        // bf16 > 0  <=>  sign bit clear and magnitude non-zero
        if (lo16 != 0u && !(lo16 & 0x8000u)) m |= 1u << (2 * j);
        if (hi16 != 0u && !(hi16 & 0x8000u)) m |= 1u << (2 * j + 1);
    }
    return m;
}
}  // namespace fz

struct FusedArgs {
    const float* x; const float* w; const float* eps;      // eps nullable (then Philox)
    const unsigned char* wimg; float* parts;                // parts: [gridDim.x][Ppad]
    float* per_jet; double* sums; float* kld_out;           // per_jet nullable [4,B]; kld_out nullable [B]
    long B; int n_tiles; int train; int mse;
    float beta;
    unsigned long long seed, row_offset; unsigned int step;
    long long* dbg;                                         // optional: clock64() stamps of CTA 0 (debug)
};

__global__ void __launch_bounds__(256, 1) fused_train_kernel(const __grid_constant__ FPlan P, FusedArgs a) {
    using namespace umma;
    using namespace fz;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t mma_bar, xbar[2], wbar[2];
    __shared__ uint32_t tmem_base_s;
    unsigned char* sm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, sub = warp & 3, half = warp >> 2;
    const int Rt = P.Rt, blk = P.blk, nl = P.nl, n = P.n_hidden, D = P.D, L = P.L, Lp = P.Lp;
    const int ks_jets = Rt >> 4;

    auto tile_hi = [&](int k) { return sm + P.tile_off[k]; };
    auto tile_lo = [&](int k) { return sm + P.tile_off[k] + P.tile_nblk[k] * blk; };
    auto wide_hi = [&](int s) { return sm + P.wide_off + s * 2 * blk; };
    auto wslot = [&](int s) { return sm + P.wring_off + s * P.wslot_bytes; };
    float* s_mu = reinterpret_cast<float*>(sm + P.scratch_off);
    float* s_lv = s_mu + Rt * L; float* s_eps = s_lv + Rt * L; float* s_rec = s_eps + Rt * L;   // s_rec: [2][Rt]

    // ---- one-time setup: zero the activation + wide region (pad columns must be exact zeros), barriers, TMEM
    for (int i = tid * 16; i < P.wring_off; i += 256 * 16) *reinterpret_cast<uint4*>(sm + i) = make_uint4(0, 0, 0, 0);
    if (tid == 0) { mbar_init(&mma_bar, 1); mbar_init(&xbar[0], 1); mbar_init(&xbar[1], 1); mbar_init(&wbar[0], 1); mbar_init(&wbar[1], 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    const uint32_t t_lane = tm + ((uint32_t)(32 * sub) << 16);
    const int n_ops = a.train ? P.n_wop : (P.nX + (nl - 2) + P.nY);
    uint32_t phm = 0, phx[2] = {0, 0};
    long gk = 0;                                             // weight ops consumed so far by this CTA

    auto w_issue = [&](int k, long g) {                      // thread 0: start the bulk copy of weight op k into slot g&1
        const int s = (int)(g & 1);
        mbar_expect_tx(&wbar[s], P.wop_bytes[k]);
        bulk_g2s(wslot(s), a.wimg + P.wop_off[k], P.wop_bytes[k], &wbar[s]);
    };
    // thread 0, at the start of the section that consumes op k (per-tile index): prefetch the next op, wait for this one
    auto w_begin = [&](int k, bool more_tiles) -> unsigned char* {
        if (k + 1 < n_ops) w_issue(k + 1, gk + 1);
        else if (more_tiles) w_issue(0, gk + 1);
        mbar_wait(&wbar[gk & 1], (uint32_t)((gk >> 1) & 1));
        return wslot((int)(gk & 1));
    };
    if (tid == 0 && (int)blockIdx.x < a.n_tiles) w_issue(0, 0);

    int n_dbg = 0;
    auto stamp = [&]() { if (a.dbg && blockIdx.x == 0 && tid == 0 && n_dbg < 255) a.dbg[1 + n_dbg++] = clock64(); };
    stamp();
    int tile_iter = 0;
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++tile_iter) {
        const bool more_tiles = tile + (int)gridDim.x < a.n_tiles;
        const long row0 = (long)tile * Rt;
        const int nvalid = (int)min((long)Rt, a.B - row0);
        const int row = 16 * sub + lane;                      // jet row owned in M=64 accumulators (lane < 16)
        const bool row_thr = lane < 16 && row < Rt;
        const float w_row = (row_thr && row < nvalid) ? a.w[row0 + row] : 0.f;
        float* part = a.parts + (long)blockIdx.x * P.Ppad;
        float rec_acc = 0.f, kld_r = 0.f;
        int kop = 0;                                          // per-tile weight-op index

        auto stage_x = [&](int c, int slot) {                 // [X,1] columns [64c, 64c+64) -> wide slot (hi, lo)
            unsigned char* hi = wide_hi(slot); unsigned char* lo = hi + blk;
            for (int idx = tid; idx < Rt * 16; idx += 256) {
                int r = idx >> 4, c4 = (idx & 15) << 2, col = 64 * c + c4;
                float v[4] = {0.f, 0.f, 0.f, 0.f};
                if (r < nvalid) {
                    const float* p = a.x + (row0 + r) * D + col;
                    if (col + 3 < D && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
                        float4 q = __ldg(reinterpret_cast<const float4*>(p));
                        v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) v[j] = (col + j < D) ? __ldg(p + j) : (col + j == D ? 1.f : 0.f);
                    }
                }
                __nv_bfloat16 h[4], l[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) split_bf16(v[j], h[j], l[j]);
                uint32_t off = sw128_off(r, c4);
                *reinterpret_cast<uint2*>(hi + off) = make_uint2(pack2(h[0], h[1]), pack2(h[2], h[3]));
                *reinterpret_cast<uint2*>(lo + off) = make_uint2(pack2(l[0], l[1]), pack2(l[2], l[3]));
            }
        };
        auto wait_mma = [&]() { mbar_wait(&mma_bar, phm); phm ^= 1u; tc_fence_after(); };
        auto wait_x = [&](int s) { mbar_wait(&xbar[s], phx[s]); phx[s] ^= 1u; tc_fence_after(); };
        auto publish = [&]() { fence_async_smem(); tc_fence_before(); __syncthreads(); };   // smem writes + TMEM reads done

        // epilogue: accumulator columns -> ReLU -> output tile k (+ ones column)
        auto epi_act = [&](int f) {
            const FLayerDev& ly = P.layer[f];
            unsigned char* th = tile_hi(f); unsigned char* tl = tile_lo(f);
            const int ng = (ly.out + 1 + 15) >> 4;
            for (int g = half; g < ng; g += 2) {
                float v[16];
                tmem_ld16(t_lane + ACC_F + 16 * g, v);
                if (row_thr) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        int col = 16 * g + j;
                        v[j] = col < ly.out ? (ly.relu ? fmaxf(v[j], 0.f) : v[j]) : (col == ly.out ? 1.f : 0.f);
                    }
                    store16(th, tl, blk, row, 16 * g, v);
                }
            }
        };
        // dW accumulator (lane = input feature) -> this CTA's partial gradient slice
        auto epi_dw = [&](int f, int M, int frow0, int col0, int ncols) {
            const FLayerDev& ly = P.layer[f];
            const int r_loc = (M == 128) ? 32 * sub + lane : 16 * sub + lane;
            const bool act = (M == 128 || lane < 16) && (frow0 + r_loc) <= ly.in;
            const long frow = frow0 + r_loc;
            const int ng = (ncols + 15) >> 4;
            for (int g = half; g < ng; g += 2) {
                float v[16];
                tmem_ld16(t_lane + ACC_DW + 16 * g, v);
                if (act) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        int col = col0 + 16 * g + j;
#pragma unroll
                        for (int sgi = 0; sgi < 2; ++sgi) {
                            int cc = col - ly.seg_col0[sgi];
                            if (cc >= 0 && cc < ly.seg_n[sgi] && 16 * g + j < ncols) {
                                float* dst = part + ly.seg_dst[sgi] + frow * ly.seg_ld[sgi] + cc;
                                *dst = tile_iter ? *dst + v[j] : v[j];
                            }
                        }
                    }
                }
            }
        };
        // dX accumulator -> ReLU mask of the layer input -> gradient tile written over the input activation tile
        auto epi_grad = [&](int f) {
            const FLayerDev& ly = P.layer[f];
            unsigned char* th = tile_hi(f - 1); unsigned char* tl = tile_lo(f - 1);
            const int ng = ly.inpad >> 4;
            for (int g = half; g < ng; g += 2) {
                float v[16];
                tmem_ld16(t_lane + ACC_F + 16 * g, v);
                if (row_thr) {
                    uint32_t m = positive16(th, blk, row, 16 * g);
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = ((m >> j) & 1u) && (16 * g + j < ly.in) ? v[j] : 0.f;
                    store16(th, tl, blk, row, 16 * g, v);
                }
            }
        };

        // =========================================================== forward, layer 0: stream [X,1] in 64-col chunks
        {
            const FLayerDev& ly = P.layer[0];
            const uint32_t idesc = idesc_bf16_major(64, ly.npad, 0, 0);
            for (int c = 0; c < P.nX; ++c) {
                const int s = c & 1;
                stage_x(c, s);
                publish();
                if (c > 0) wait_x(s ^ 1);                                      // MMA of chunk c-1 complete
                if (tid == 0) {
                    unsigned char* wb = w_begin(kop, more_tiles);
                    tc_fence_after();
                    const int ncol = min(64, D + 1 - 64 * c);
                    issue_kk(tm + ACC_F, smem_u32(wide_hi(s)), blk, 0, smem_u32(wb), ly.npad * 128, 0, 1, (ncol + 15) >> 4, idesc, c > 0);
                    mma_commit(&xbar[s]);
                }
                ++kop; ++gk;
            }
            wait_x((P.nX - 1) & 1);
            stamp();
            epi_act(0);
            publish();
            stamp();
        }
        // =========================================================== forward chain: layers 1 .. nl-2
        for (int f = 1; f <= nl - 2; ++f) {
            const FLayerDev& ly = P.layer[f];
            if (tid == 0) {
                unsigned char* wb = w_begin(kop, more_tiles);
                tc_fence_after();
                const int last = ly.in + 1 - 64 * (ly.kb_fwd - 1);
                issue_kk(tm + ACC_F, smem_u32(tile_hi(f - 1)), P.tile_nblk[f - 1] * blk, blk, smem_u32(wb), ly.kb_fwd * ly.npad * 128,
                         ly.npad * 128, ly.kb_fwd, (last + 15) >> 4, idesc_bf16_major(64, ly.npad, 0, 0), false);
                mma_commit(&mma_bar);
            }
            ++kop; ++gk;
            wait_mma();
            if (f != n) {
                epi_act(f);
            } else if (half == 0) {
                // ---- heads: Sampling (models.py:9-15) + KLD (models.py:81-85); z -> tile n
                float kl = 0.f;
                for (int q = 0; q < (L + 15) >> 4; ++q) {
                    float mu[16], lv[16], z[16];
                    tmem_ld16(t_lane + ACC_F + 16 * q, mu);
                    tmem_ld16(t_lane + ACC_F + Lp + 16 * q, lv);
                    if (row_thr) {
                        float nrm[4];
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            int l = 16 * q + j;
                            z[j] = 0.f;
                            if (l < L) {
                                float e;
                                if (a.eps) e = row < nvalid ? a.eps[(row0 + row) * L + l] : 0.f;
                                else {
                                    if ((l & 3) == 0) philox_normal4(a.seed, a.row_offset + (unsigned long long)(row0 + row), (unsigned)(l >> 2), a.step, nrm);
                                    e = nrm[l & 3];
                                }
                                float ex = clipv(expf(lv[j]));
                                kl += 1.f + lv[j] - ex - mu[j] * mu[j];
                                float sg = clipv(expf(lv[j] * 0.5f));
                                z[j] = mu[j] + sg * e;
                                s_mu[row * L + l] = mu[j]; s_lv[row * L + l] = lv[j]; s_eps[row * L + l] = e;
                            } else if (l == L) z[j] = 1.f;
                        }
                        store16(tile_hi(n), tile_lo(n), blk, row, 16 * q, z);
                    }
                }
                if (L % 16 == 0 && row_thr) {                                  // ones column falls into the next 16-group
                    float z[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) z[j] = j == 0 ? 1.f : 0.f;
                    store16(tile_hi(n), tile_lo(n), blk, row, L, z);
                }
                kld_r = -(kl / (float)L) / 2.f;
            }
            publish();
            stamp();
        }
        // =========================================================== output layer in 64-column chunks
        {
            const int f = nl - 1;
            const FLayerDev& ly = P.layer[f];
            const int a_nblk = P.tile_nblk[f - 1];
            const int lastk = ly.in + 1 - 64 * (ly.kb_fwd - 1);
            const uint32_t fwd_lo = ly.kb_fwd * 64 * 128;                      // lo part of the fwd sub-blob
            const uint32_t dx_off = 2 * fwd_lo, dx_lo = ly.inpad * 128;
            for (int c = 0; c < P.nY; ++c) {
                unsigned char* wb = nullptr;
                if (tid == 0) {
                    wb = w_begin(kop, more_tiles);
                    stamp();
                    tc_fence_after();
                    issue_kk(tm + ACC_Y, smem_u32(tile_hi(f - 1)), a_nblk * blk, blk, smem_u32(wb), fwd_lo, 64 * 128, ly.kb_fwd,
                             (lastk + 15) >> 4, idesc_bf16_major(64, 64, 0, 0), false);
                    mma_commit(&mma_bar);
                }
                wait_mma();
                stamp();
                // reconstruction loss (models.py:66,72-78) and its gradient through clip_values
                unsigned char* gh = wide_hi(c & 1); unsigned char* gl = gh + blk;
                for (int gq = 0; gq < 2; ++gq) {
                    const int g = 2 * half + gq;
                    float v[16];
                    tmem_ld16(t_lane + ACC_Y + 16 * g, v);
                    if (row_thr) {
                        const int col0 = 64 * c + 16 * g;
                        float xv[16];
                        const float* px = a.x + (row0 + row) * D + col0;
                        if (row < nvalid && col0 + 15 < D && ((reinterpret_cast<uintptr_t>(px) & 15) == 0)) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                float4 q = __ldg(reinterpret_cast<const float4*>(px) + j);
                                xv[4 * j] = q.x; xv[4 * j + 1] = q.y; xv[4 * j + 2] = q.z; xv[4 * j + 3] = q.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) xv[j] = (row < nvalid && col0 + j < D) ? __ldg(px + j) : 0.f;
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            float gy = 0.f;
                            if (col0 + j < D) {
                                float yr = v[j], d = clipv(yr) - xv[j];
                                rec_acc += a.mse ? d * d : fabsf(d);
                                float gg = a.mse ? w_row * 2.f * d / (float)D : w_row * (float)((d > 0.f) - (d < 0.f)) / (float)D;
                                gy = clip_pass(yr) ? gg : 0.f;
                            }
                            v[j] = gy;
                        }
                        if (a.train) store16(gh, gl, blk, row, 16 * g, v);
                    }
                }
                stamp();
                publish();
                stamp();
                if (a.train) {
                    if (tid == 0) {
                        tc_fence_after();
                        // dW_out[:, chunk] = [d_n,1]^T g_y   (MN-major A and B, K = jets)
                        const int M = a_nblk >= 2 ? 128 : 64;
                        issue_mn(tm + ACC_DW, smem_u32(tile_hi(f - 1)), a_nblk * blk, smem_u32(gh), blk, blk, ks_jets, idesc_bf16_major(M, 64, 1, 1));
                        // g_dn += g_y[:, chunk] * W_out[:, chunk]^T
                        const int ncol = min(64, D - 64 * c);
                        issue_kk(tm + ACC_F, smem_u32(gh), blk, 0, smem_u32(wb) + dx_off, dx_lo, 0, 1, (ncol + 15) >> 4,
                                 idesc_bf16_major(64, ly.inpad, 0, 0), c > 0);
                        mma_commit(&mma_bar);
                    }
                    wait_mma();
                    stamp();
                    epi_dw(f, a_nblk >= 2 ? 128 : 64, 0, 64 * c, min(64, D - 64 * c));
                    stamp();
                    publish();
                }
                ++kop; ++gk;
                stamp();
            }
            if (a.train) { epi_grad(f); publish(); }
            stamp();
        }
        // =========================================================== backward chain: layers nl-2 .. 1
        if (a.train) {
            for (int f = nl - 2; f >= 1; --f) {
                const FLayerDev& ly = P.layer[f];
                const int a_nblk = P.tile_nblk[f - 1];
                const int M = a_nblk >= 2 ? 128 : 64;
                if (tid == 0) {
                    unsigned char* wb = w_begin(kop, more_tiles);
                    tc_fence_after();
                    issue_mn(tm + ACC_DW, smem_u32(tile_hi(f - 1)), a_nblk * blk, smem_u32(tile_hi(f)), P.tile_nblk[f] * blk, blk, ks_jets,
                             idesc_bf16_major(M, ly.npad, 1, 1));
                    const int lastk = ly.out - 64 * (ly.kb_dx - 1);
                    issue_kk(tm + ACC_F, smem_u32(tile_hi(f)), P.tile_nblk[f] * blk, blk, smem_u32(wb), ly.kb_dx * ly.inpad * 128,
                             ly.inpad * 128, ly.kb_dx, (lastk + 15) >> 4, idesc_bf16_major(64, ly.inpad, 0, 0), false);
                    mma_commit(&mma_bar);
                }
                ++kop; ++gk;
                wait_mma();
                epi_dw(f, M, 0, 0, ly.npad);
                if (f != n + 1) {
                    epi_grad(f);
                } else if (half == 0) {
                    // ---- latent backward: g_z -> (g_mu | g_lv) written over the z tile
                    const float gk_row = a.beta * w_row;
                    for (int q = 0; q < (2 * Lp) >> 4; ++q) {
                        float gz[16], o[16];
                        const bool is_lv = 16 * q >= Lp;
                        tmem_ld16(t_lane + ACC_F + (is_lv ? 16 * q - Lp : 16 * q), gz);
                        if (row_thr) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                int l = (is_lv ? 16 * q - Lp : 16 * q) + j;
                                o[j] = 0.f;
                                if (l < L) {
                                    float mu = s_mu[row * L + l], lv = s_lv[row * L + l];
                                    if (!is_lv) o[j] = gz[j] + gk_row * mu / (float)L;
                                    else {
                                        float e_raw = expf(lv), s_raw = expf(lv * 0.5f);
                                        o[j] = gk_row * (-(1.f - (clip_pass(e_raw) ? e_raw : 0.f)) / (2.f * (float)L))
                                             + gz[j] * s_eps[row * L + l] * (clip_pass(s_raw) ? s_raw * 0.5f : 0.f);
                                    }
                                }
                            }
                            store16(tile_hi(n), tile_lo(n), blk, row, 16 * q, o);
                        }
                    }
                }
                publish();
                stamp();
            }
            // ======================================================= layer 0: dW = [X,1]^T g_0, X re-streamed (L2)
            {
                const FLayerDev& ly = P.layer[0];
                for (int c = 0; c < P.nX; ++c) {
                    stage_x(c, c & 1);
                    publish();
                    if (tid == 0) {
                        tc_fence_after();
                        issue_mn(tm + ACC_DW, smem_u32(wide_hi(c & 1)), blk, smem_u32(tile_hi(0)), P.tile_nblk[0] * blk, blk, ks_jets,
                                 idesc_bf16_major(64, ly.npad, 1, 1));
                        mma_commit(&mma_bar);
                    }
                    wait_mma();
                    epi_dw(0, 64, 64 * c, 0, ly.npad);
                    publish();
                }
            }
        }
        stamp();
        // =========================================================== per-jet losses (models.py:123-132) and batch sums
        if (row_thr) s_rec[half * Rt + row] = rec_acc;
        __syncthreads();
        {
            float lr_ = 0.f, lk = 0.f, tot = 0.f;
            if (half == 0 && row_thr && row < nvalid) {
                float rec = (s_rec[row] + s_rec[Rt + row]) / (float)D;
                lr_ = rec * w_row; lk = kld_r * w_row * a.beta; tot = lr_ + lk + 0.f;
                if (a.per_jet) {
                    a.per_jet[row0 + row] = lr_; a.per_jet[a.B + row0 + row] = lk;
                    a.per_jet[2 * a.B + row0 + row] = 0.f; a.per_jet[3 * a.B + row0 + row] = tot;
                }
                if (a.kld_out) a.kld_out[row0 + row] = kld_r;
            }
            if (a.sums && half == 0) {
                double v0 = warp_sum((double)lr_), v1 = warp_sum((double)lk), v3 = warp_sum((double)tot);
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

// ------------------------------------------------------------------------------------------------
// Sum of the per-CTA partial gradients (fixed order), clip_values (models.py:185), Keras Adam (models.py:186) and
// refresh of the bf16 hi/lo weight images consumed by fused_train_kernel.
// ------------------------------------------------------------------------------------------------
struct AdamImgArgs {
    const float* parts; int n_parts; long part_stride;
    const float* grad_in; float* grad_out;
    float* theta; float* m; float* v; long n;
    float alpha, b1, b2, eps; int apply;
    unsigned char* wimg; const int* pos;                    // pos: [4][n] byte offsets (fwd hi, fwd lo, dx hi, dx lo), -1 = none
};

// block = (64 parameters) x (8 slices of the partial-gradient list): slice y sums parts y, y+8, ... in order, ten
// independent loads in flight at a time (one L2 round trip for the 79 slices of a B=5000 step: the kernel is bound by
// L2 latency, not bandwidth), one resident wave of 512-thread blocks; the eight slice sums are
// combined in a fixed tree, so the result does not depend on scheduling.
constexpr int kAdamCols = 64, kAdamSlices = 8;
__global__ void __launch_bounds__(kAdamCols * kAdamSlices, 4) adam_img_kernel(AdamImgArgs a) {
    __shared__ float red[kAdamSlices][kAdamCols + 1];
    // programmatic dependent launch (no-ops otherwise): let the next training kernel begin its set-up, then wait for
    // the kernel that produced the partial gradients
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int tx = threadIdx.x & (kAdamCols - 1), ty = threadIdx.x / kAdamCols;
    const long i = (long)blockIdx.x * kAdamCols + tx;
    const bool in = i < a.n;
    float g = 0.f;
    if (a.parts) {
        if (in) {
            const float* src = a.parts + i;
            for (int base = ty; base < a.n_parts; base += 10 * kAdamSlices) {
                float t[10];
#pragma unroll
                for (int k = 0; k < 10; ++k) {
                    const int sidx = base + kAdamSlices * k;
                    t[k] = sidx < a.n_parts ? __ldcg(src + (long)sidx * a.part_stride) : 0.f;
                }
                g += (((t[0] + t[1]) + (t[2] + t[3])) + ((t[4] + t[5]) + (t[6] + t[7]))) + (t[8] + t[9]);
            }
        }
        red[ty][tx] = g;
        __syncthreads();
        if (ty == 0) g = ((red[0][tx] + red[1][tx]) + (red[2][tx] + red[3][tx])) + ((red[4][tx] + red[5][tx]) + (red[6][tx] + red[7][tx]));
    }
    if (ty != 0 || !in) return;
    float th = a.theta[i];
    if (a.parts || a.grad_in) {
        if (a.parts) { if (a.grad_out) a.grad_out[i] = g; }
        else g = a.grad_in[i];
        if (a.apply) {
            g = clipv(g);
            float m = a.m[i], v = a.v[i];
            m += (g - m) * (1.f - a.b1);
            v += (g * g - v) * (1.f - a.b2);
            a.m[i] = m; a.v[i] = v;
            th -= (m * a.alpha) / (sqrtf(v) + a.eps);
            a.theta[i] = th;
        }
    }
    if (a.wimg && (a.apply || !(a.parts || a.grad_in))) {
        __nv_bfloat16 hi, lo;
        umma::split_bf16(th, hi, lo);
        int p0 = a.pos[i], p1 = a.pos[a.n + i], p2 = a.pos[2 * a.n + i], p3 = a.pos[3 * a.n + i];
        if (p0 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p0) = hi;
        if (p1 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p1) = lo;
        if (p2 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p2) = hi;
        if (p3 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p3) = lo;
    }
}

}  // namespace oevae
