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

constexpr uint32_t T_MU = 0, T_LV = 64, T_CH = 128, T_Y0 = 192, T_DW = 320, T_C2 = 448;   // T_C2: second chain accumulator

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
// D[128 x N] (+)= A * B over nk k-steps, bf16x3: Ahi*Bhi + Ahi*Blo + Alo*Bhi.  Descriptors are built once; a k-step
// only advances their 14-bit start-address field (shared memory is < 256 KB, so the field cannot overflow).
__device__ __noinline__ void issue_core(uint32_t d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t a_step, uint32_t b_step,
                                        int nk, uint32_t idesc, uint32_t acc) {
    for (int ks = 0; ks < nk; ++ks) {
        mma_bf16_ss(d, ah, bh, idesc, acc);
        mma_bf16_ss(d, ah, bl, idesc, 1u);
        mma_bf16_ss(d, al, bh, idesc, 1u);
        acc = 1u;
        ah += a_step; al += a_step; bh += b_step; bl += b_step;
    }
}
// same, k-steps alternating between two accumulators (d0 even, d1 odd) so that consecutive MMAs are independent;
// acc0 / acc1: whether d0 / d1 already hold partial sums.  The epilogue adds the two.
__device__ __noinline__ void issue_core2(uint32_t d0, uint32_t d1, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl, uint32_t a_step,
                                         uint32_t b_step, int nk, uint32_t idesc, uint32_t acc0, uint32_t acc1) {
    for (int ks = 0; ks < nk; ++ks) {
        const uint32_t d = (ks & 1) ? d1 : d0, acc = (ks & 1) ? acc1 : acc0;
        mma_bf16_ss(d, ah, bh, idesc, acc);
        mma_bf16_ss(d, ah, bl, idesc, 1u);
        mma_bf16_ss(d, al, bh, idesc, 1u);
        if (ks & 1) acc1 = 1u; else acc0 = 1u;
        ah += a_step; al += a_step; bh += b_step; bl += b_step;
    }
}
__device__ __forceinline__ void issue2(uint32_t d0, uint32_t d1, const Opnd& A, const Opnd& B, int nk, int N, bool acc0, bool acc1) {
    issue_core2(d0, d1, desc(A, 0, false), desc(A, 0, true), desc(B, 0, false), desc(B, 0, true), A.kstep >> 4, B.kstep >> 4, nk,
                idesc_bf16_major(128, N, A.mn, B.mn), acc0 ? 1u : 0u, acc1 ? 1u : 0u);
}
__device__ __forceinline__ void issue(uint32_t d, const Opnd& A, const Opnd& B, int nk, int N, bool acc) {
    issue_core(d, desc(A, 0, false), desc(A, 0, true), desc(B, 0, false), desc(B, 0, true), A.kstep >> 4, B.kstep >> 4, nk,
               idesc_bf16_major(128, N, A.mn, B.mn), acc ? 1u : 0u);
}

// x0, x1 -> packed bf16 hi parts and packed bf16 lo parts (x ~= hi + lo, |err| <= 2^-17 |x|); non-finite residuals
// (x = +-inf) are dropped so that an overflowing activation stays +-inf instead of turning into NaN
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    const __nv_bfloat162 hh = __floats2bfloat162_rn(x0, x1);
    hi = *reinterpret_cast<const uint32_t*>(&hh);
    float r0 = x0 - __uint_as_float(hi << 16), r1 = x1 - __uint_as_float(hi & 0xffff0000u);
    r0 = (fabsf(x0) <= 3.0e38f) ? r0 : 0.f; r1 = (fabsf(x1) <= 3.0e38f) ? r1 : 0.f;
    const __nv_bfloat162 ll = __floats2bfloat162_rn(r0, r1);
    lo = *reinterpret_cast<const uint32_t*>(&ll);
}
// 16 consecutive jets (col0 % 16 == 0) of feature row `row` of an activation tile, as bf16 hi/lo
__device__ __forceinline__ void store16(unsigned char* t_hi, unsigned char* t_lo, int row, int col0, const float* v) {
    uint32_t h[8], l[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) split2(v[2 * j], v[2 * j + 1], h[j], l[j]);
    const uint32_t o0 = sw128_off(row, col0), o1 = sw128_off(row, col0 + 8);
    *reinterpret_cast<uint4*>(t_hi + o0) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(t_hi + o1) = make_uint4(h[4], h[5], h[6], h[7]);
    *reinterpret_cast<uint4*>(t_lo + o0) = make_uint4(l[0], l[1], l[2], l[3]);
    *reinterpret_cast<uint4*>(t_lo + o1) = make_uint4(l[4], l[5], l[6], l[7]);
}
// 8 consecutive jets (col0 % 8 == 0) of feature row `row`
__device__ __forceinline__ void store8(unsigned char* t_hi, unsigned char* t_lo, int row, int col0, const float* v) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) split2(v[2 * j], v[2 * j + 1], h[j], l[j]);
    const uint32_t o = sw128_off(row, col0);
    *reinterpret_cast<uint4*>(t_hi + o) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(t_lo + o) = make_uint4(l[0], l[1], l[2], l[3]);
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
// 16 values per lane summed over the 32 lanes; every lane returns the total of index (lane & 15)
__device__ __forceinline__ float transpose_reduce16(float* v, int lane) {
#pragma unroll
    for (int s = 8; s >= 1; s >>= 1) {
        const bool up = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const float send = up ? v[i] : v[i + s], keep = up ? v[i + s] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}
// 8 values per lane summed over the 16 lanes of a half warp; every lane returns the total of index (lane & 7)
__device__ __forceinline__ float transpose_reduce8(float* v, int lane) {
#pragma unroll
    for (int s = 4; s >= 1; s >>= 1) {
        const bool up = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const float send = up ? v[i] : v[i + s], keep = up ? v[i + s] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 8);
}
// nan_to_num(p * log2(p / q)) of utils.py:347-350 on the SFU: approximate reciprocal + lg2.approx (abs error of the log
// <= 2^-22 near 1, 2 ulp elsewhere); p/q < 0 -> NaN -> 0, q = 0 -> +-inf -> +-FLT_MAX or NaN -> 0 exactly as NumPy's
__device__ __forceinline__ float kld_term_fast(float p, float q) {
    const float t = p * __log2f(__fdividef(p, q));
    return (t != t) ? 0.f : fminf(fmaxf(t, -3.402823466e+38f), 3.402823466e+38f);
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
    // SCORE instantiation only (anomaly scoring, vae.py:163-167 + utils.py:346-386): nullable [B] outputs
    float* s_mse; float* s_mae; float* s_kld; float* s_jsd; float* s_lat; unsigned char* s_fin;
    float* s_xhat;                                          // nullable [B, D]: clip_values(X_hat) of the pass (vae.py:165)
    int* tile_ctr;                                          // fused3 SCORE: zeroed before the launch; tiles beyond the first gridDim.x are handed out from it
    int heads_t;                                            // fused3 SCORE: heads with the jets on the TMEM lanes (M = 64 accumulators)
    // fused3 only -- outlier-exposure step (models.py:86-132): rows [B1, B) of the batch are read from x2 (the OoD sample) and
    // start at a tile boundary (n_tiles = ceil(B1 / 64) + ceil((B - B1) / 64)); the loss-term coefficients come per jet from
    // c_rec / c_kld (nullable pair; else w and beta w, or 1 and 1 when w is null too); row i of either half is given the noise
    // of row i, as the reference's fixed-seed Sampling layer gives them
    const float* x2; long B1; const float* c_rec; const float* c_kld;
    // paired mode (OE_type 'KLD'): B = number of (bkg, OoD) pairs, tile t = pairs [32 t, 32 t + 32) with the bkg rows (x, w) on jet
    // columns 0..31 and their OoD partners (x2, w2) on columns 32..63; per_jet / sums get all four loss rows of models.py:123-132
    int paired; const float* w2; float lamb, margin;
    // fused3 training only, nullable: two monotonic counters.  Every drain warp of a CTA adds 1 to grp_ctr[0] once the CTA's slice of the
    // OUTPUT layer's gradient is final (after the output chunks of its last tile) and 1 to grp_ctr[1] once the chain layers' gradients are
    // (everything but layer 0).  The update kernel's blocks for those parameters start on them while the rest of the backward pass runs
    // (adam_kernels.cuh: EarlyGroups).
    unsigned int* grp_ctr;
};


constexpr int kF2Threads = 512;                             // 16 warps: lane quarter q = warp & 3, column group h = warp >> 2

// SCORE = false: training step (a.train) / get_losses (eval).  SCORE = true: forward only, and the output epilogue
// accumulates loss_function's MSE / MAE / KLD / JSD between X and clip_values(X_hat) instead of the training loss.
template <bool SCORE>
__global__ void __launch_bounds__(kF2Threads, 1) fused2_kernel(const __grid_constant__ F2Plan P, F2Args a) {
    const bool is_train = !SCORE && a.train;
    using namespace umma;
    using namespace f2;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t mma_bar, y_bar[2], xbar[kF2XSlots], wbar[kF2WSlots];
    __shared__ uint32_t tmem_base_s;
    unsigned char* sm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = warp & 3, h = warp >> 2;
    const int frow = 32 * q + lane;                          // TMEM lane = feature row owned by this thread
    const int jc0 = 16 * h;                                  // first of the 16 jet columns owned by this thread
    const int n = P.n, D = P.D, L = P.L, so = 2 * n + 1;
    const int nops = is_train ? P.n_op_train : P.n_op_eval;
    const int n_fwd_ops = P.op_first[so] + P.nX;

    float* s_w = reinterpret_cast<float*>(sm + P.scratch_off);        // [64] jet weights (0 for rows past the batch)
    float* s_rec = s_w + 64;                                           // [4][64] per lane-quarter reconstruction sums
    float* s_kld = s_rec + 256;                                        // [64]
    float* s_db = s_kld + 64;                                          // [4 buffers][4 column groups][128] bias-gradient parts
    float* s_eps = s_db + 2048;                                        // [L][64]
    auto tile_hi = [&](int k) { return sm + P.tile_off[k]; };
    auto tile_lo = [&](int k) { return sm + P.tile_off[k] + P.tile_rows[k] * 128; };
    auto xslot = [&](int s) { return sm + P.s_off + s * 16384; };      // hi 8 KB, lo 8 KB
    unsigned char* gy_hi = sm + P.s_off; unsigned char* gy_lo = gy_hi + 16384;
    auto wslot = [&](int g) { return sm + P.wring_off + (g & (kF2WSlots - 1)) * P.wslot_bytes; };
    auto db_at = [&](int buf, int hh, int row) -> float& { return s_db[(buf * 4 + hh) * 128 + row]; };

    // ---- one-time setup
    if (tid == 0) {
        mbar_init(&mma_bar, 1); mbar_init(&y_bar[0], 1); mbar_init(&y_bar[1], 1);
        for (int i = 0; i < kF2XSlots; ++i) mbar_init(&xbar[i], 1);
        for (int i = 0; i < kF2WSlots; ++i) mbar_init(&wbar[i], 1);
        mbar_fence_init();
    }
    for (int i = tid * 16; i < P.zero_bytes; i += kF2Threads * 16) *reinterpret_cast<uint4*>(sm + i) = make_uint4(0, 0, 0, 0);
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    const uint32_t t_lane = tm + ((uint32_t)(32 * q) << 16);

    // ---- weight ring: ops are consumed strictly in order; op g lives in slot g % 4.  Thread 0 (the MMA issuer) waits for
    // arrivals (wget); the producer (lane 0 of the last warp, idle in most epilogues) mirrors the same program order,
    // declares ops consumed (wdone) and keeps the ring full.  Each keeps its own counters; nothing is shared.
    const bool producer = tid == kF2Threads - 32;
    const int my_tiles = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const int w_total = my_tiles * nops;
    int w_issued = 0, w_done = 0, gbase = 0, w_next = 0;
    auto wbar_wait = [&](int g) { mbar_wait(&wbar[g & (kF2WSlots - 1)], (uint32_t)((g >> 2) & 1)); };
    auto refill = [&]() {
        while (w_issued < w_total && w_issued < w_done + kF2WSlots) {
            const int k = w_issued % nops, s = w_issued & (kF2WSlots - 1);
            const uint32_t bytes = (!is_train && k >= n_fwd_ops) ? 16u : P.op_bytes[k];
            mbar_expect_tx(&wbar[s], bytes);
            bulk_g2s(wslot(w_issued), a.wimg + P.op_off[k], bytes, &wbar[s]);
            ++w_issued;
        }
    };
    auto wget = [&](int k) -> uint32_t {                      // thread 0: wait until op k of this tile has landed; returns its slot
        for (; w_next <= k; ++w_next) wbar_wait(gbase + w_next);
        return smem_u32(wslot(gbase + k));
    };
    auto wdone = [&](int k) { w_done = gbase + k + 1; refill(); };   // producer: every MMA that reads ops <= k has completed
    auto wdone_dummy = [&](int k) { wbar_wait(gbase + k); wdone(k); };   // producer: op k has no consumer, only its arrival counts
    if (producer) refill();

    uint32_t ph_m = 0, ph_y = 0, ph_x = 0;
    auto wait_mma = [&]() { mbar_wait(&mma_bar, ph_m); ph_m ^= 1u; tc_fence_after(); };
    // one barrier per accumulator slot: y(c+1) is issued by thread 0 right after ITS wait on y(c), so a single barrier could
    // complete twice before a slow warp has polled it once (parity aliasing)
    auto wait_y = [&](int c) { mbar_wait(&y_bar[c & 1], (ph_y >> (c & 1)) & 1u); ph_y ^= 1u << (c & 1); tc_fence_after(); };
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
        if (tid < 64) s_w[tid] = (tid < nvalid && !SCORE) ? a.w[row0 + tid] : 0.f;
        for (int pf = tile_iter ? 1 : 0; pf < 2; ++pf) {   // pull the next tile's rows of X (and the first tile's own) towards L2
            const long nrow0 = row0 + (long)pf * gridDim.x * kF2Jets;
            const long nbytes = min((long)kF2Jets, a.B - nrow0) * D * 4;
            const char* nx = reinterpret_cast<const char*>(a.x + nrow0 * D);
            for (long o = (long)tid * 128; o < nbytes; o += kF2Threads * 128) prefetch_l2(nx + o);
        }
        // reparameterisation noise of the tile -> s_eps[l][jet]: external eps, or one Philox counter (row, quad, step)
        // = latents 4*quad .. 4*quad+3 of one jet (the spec of oracle/philox.py)
        if (a.eps) {
            for (int i = tid; i < 64 * L; i += kF2Threads) {
                const int j = i / L, l = i - j * L;
                s_eps[l * 64 + j] = j < nvalid ? __ldg(a.eps + (row0 + j) * L + l) : 0.f;
            }
        } else if (tid < 64 * ((L + 3) >> 2)) {
            const int j = tid & 63, quad = tid >> 6;
            float nrm[4];
            philox_normal4(a.seed, a.row_offset + (unsigned long long)(row0 + j), (unsigned)quad, a.step, nrm);
#pragma unroll
            for (int t = 0; t < 4; ++t) if (4 * quad + t < L) s_eps[(4 * quad + t) * 64 + j] = nrm[t];
        }

        // X chunk c = [64 jets x features 64c..64c+63]: ldx pulls this thread's two float4 into registers (issued two
        // chunks ahead), stx writes them to ring slot xs as bf16 hi/lo (K-major rows = jets)
        auto ldx = [&](int c, float4 (&r)[2]) {
#pragma unroll
            for (int it = 0; it < 2; ++it) {
                const int idx = tid + kF2Threads * it, rr = idx >> 4, c4 = (idx & 15) << 2, col = 64 * c + c4;
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                if (c < P.nX && rr < nvalid) {
                    const float* p = a.x + (row0 + rr) * D + col;
                    if (col + 3 < D && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) t = __ldg(reinterpret_cast<const float4*>(p));
                    else {
                        if (col < D) t.x = __ldg(p);
                        if (col + 1 < D) t.y = __ldg(p + 1);
                        if (col + 2 < D) t.z = __ldg(p + 2);
                        if (col + 3 < D) t.w = __ldg(p + 3);
                    }
                }
                r[it] = t;
            }
        };
        auto stx = [&](int xs, const float4 (&r)[2]) {
            unsigned char* hi = xslot(xs); unsigned char* lo = hi + 8192;
#pragma unroll
            for (int it = 0; it < 2; ++it) {
                const int idx = tid + kF2Threads * it, rr = idx >> 4, c4 = (idx & 15) << 2;
                uint32_t ha, la, hb, lb;
                split2(r[it].x, r[it].y, ha, la); split2(r[it].z, r[it].w, hb, lb);
                const uint32_t off = sw128_off(rr, c4);
                *reinterpret_cast<uint2*>(hi + off) = make_uint2(ha, hb);
                *reinterpret_cast<uint2*>(lo + off) = make_uint2(la, lb);
            }
        };
        // chain accumulator -> bias + ReLU -> tile s (pad rows zero)
        auto epi_act = [&](int s, float bias, bool two) {
            const F2Stage& st = P.st[s];
            if (32 * q >= P.tile_rows[s]) return;
            float v[16];
            tmem_ld16(t_lane + T_CH + jc0, v);
            if (two) {
                float u[16];
                tmem_ld16(t_lane + T_C2 + jc0, u);
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] += u[j];
            }
            if (frow < P.tile_rows[s]) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float t = v[j] + bias;
                    v[j] = frow < st.out ? (st.relu ? fmaxf(t, 0.f) : t) : 0.f;
                }
                store16(tile_hi(s), tile_lo(s), frow, jc0, v);
            }
        };
        auto bias_of = [&](int s) { return frow < P.st[s].out ? __ldg(a.theta + P.st[s].seg_b[0] + frow) : 0.f; };
        // chain accumulator (dX of stage s) -> ReLU mask of H_{s-1} -> G_{s-1} over tile s-1, bias-gradient parts
        auto epi_grad = [&](int s) {
            const F2Stage& st = P.st[s];
            const int rows = P.tile_rows[s - 1];
            if (32 * q >= rows) return;
            float v[16];
            tmem_ld16(t_lane + T_CH + jc0, v);
            if (frow < rows) {
                float dbs = 0.f;
                const uint32_t m = positive16(tile_hi(s - 1), frow, jc0);
#pragma unroll
                for (int j = 0; j < 16; ++j) { v[j] = (((m >> j) & 1u) && frow < st.in) ? v[j] : 0.f; dbs += v[j]; }
                store16(tile_hi(s - 1), tile_lo(s - 1), frow, jc0, v);
                db_at((s - 1) & 1, h, frow) = dbs;
            }
        };
        // 16 columns i0.. of a dW^T accumulator row -> rows i0.. of the Keras [in, ld] kernel gradient, column o
        auto put_dw = [&](float* dw, long ld, int i0, int in, const float* v) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                if (i0 + j < in) { float* d = dw + (long)(i0 + j) * ld; *d = tile_iter ? *d + v[j] : v[j]; }
            }
        };
        auto put_db = [&](float* d, int buf, int row) {
            const float t = (db_at(buf, 0, row) + db_at(buf, 1, row)) + (db_at(buf, 2, row) + db_at(buf, 3, row));
            *d = tile_iter ? *d + t : t;
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
            for (int g = h; g < (st.kin >> 4); g += 4) {
                float v[16];
                tmem_ld16(t_lane + T_DW + 16 * g, v);
                if (sg >= 0) put_dw(dw, ld, 16 * g, st.in, v);
            }
            if (sg >= 0 && h == 0) put_db(part + st.seg_b[sg] + o, buf, frow);
        };

        // =========================================================== forward, stage 0: X streamed in 64-feature chunks
        {
            const F2Stage& st = P.st[0];
            float4 ra[2], rb[2];
            ldx(0, ra); ldx(1, rb);
            auto chunk = [&](int c, float4 (&r)[2]) {
                const int xs = c % kF2XSlots;
                if (c >= kF2XSlots) { wait_x(xs); if (producer) wdone(P.op_first[0] + c - kF2XSlots); }
                stx(xs, r);
                ldx(c + 2, r);
                publish();
                if (tid == 0) {
                    const uint32_t wb = wget(P.op_first[0] + c);
                    tc_fence_after();
                    const int nk = min(4, (D - 64 * c + 15) >> 4);
                    issue2(tm + T_CH, tm + T_C2, opnd_k(wb, st.rows * 128), opnd_k(smem_u32(xslot(xs)), 8192), nk, kF2Jets, c > 0, c > 0);
                    mma_commit(&xbar[xs]);
                }
            };
            for (int c = 0; c < P.nX; c += 2) {
                chunk(c, ra);
                if (c + 1 < P.nX) chunk(c + 1, rb);
            }
            const float bias0 = bias_of(0);
            for (int c = max(0, P.nX - kF2XSlots); c < P.nX; ++c) wait_x(c % kF2XSlots);
            if (producer) wdone(P.op_first[0] + P.nX - 1);
            stamp();
            epi_act(0, bias0, D > 16);
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
                if (false) {}
                else if (st.kind == 1) issue(tm + T_CH, opnd_mn(wb, part_b, part_b), B, nk, kF2Jets, false);
                else if (st.kind == 2) issue(tm + T_CH, opnd_k(wb, part_b), B, nk, kF2Jets, false);
                else {                                                   // heads: mean block, log_var block
                    issue(tm + T_MU, opnd_mn(wb, 2 * part_b, part_b), B, nk, kF2Jets, false);
                    issue(tm + T_LV, opnd_mn(wb + part_b, 2 * part_b, part_b), B, nk, kF2Jets, false);
                }
                mma_commit(&mma_bar);
            }
            // biases are fetched while the MMAs run
            float b_mu = 0.f, b_lv = 0.f;
            if (s != n) b_mu = bias_of(s);
            else if (lane < L) { b_mu = __ldg(a.theta + st.seg_b[0] + lane); b_lv = __ldg(a.theta + st.seg_b[1] + lane); }
            wait_mma();
            if (producer) wdone(P.op_first[s]);
            if (s != n) {
                epi_act(s, b_mu, false);
            } else if (q == 0) {
                // ---- heads: Sampling (models.py:9-15) + KLD (models.py:81-85); z -> rows 0..15 of the latent tile
                float mu[16], lv[16], kl[16];
                tmem_ld16(t_lane + T_MU + jc0, mu);
                tmem_ld16(t_lane + T_LV + jc0, lv);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    float zz = 0.f, kk = 0.f;
                    if (lane < L) {
                        const float m = mu[j] + b_mu, v = lv[j] + b_lv;
                        zz = m + clipv(expf(v * 0.5f)) * s_eps[lane * 64 + jc0 + j];
                        kk = 1.f + v - clipv(expf(v)) - m * m;
                    }
                    mu[j] = zz; kl[j] = kk;
                }
                if (lane < 16) store16(tile_hi(n), tile_lo(n), lane, jc0, mu);
                const float ks = transpose_reduce16(kl, lane);
                if (lane < 16) s_kld[jc0 + lane] = -(ks / (float)L) / 2.f;
            }
            publish();
            stamp();
        }
        // =========================================================== output layer in 128-feature chunks
        float rec_acc[16];
        float sc_mae[SCORE ? 16 : 1], sc_kl[SCORE ? 16 : 1], sc_js[SCORE ? 16 : 1];   // rec_acc doubles as the MSE sum
#pragma unroll
        for (int j = 0; j < 16; ++j) rec_acc[j] = 0.f;
        if constexpr (SCORE) {
#pragma unroll
            for (int j = 0; j < 16; ++j) { sc_mae[j] = 0.f; sc_kl[j] = 0.f; sc_js[j] = 0.f; }
        }
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
                mma_commit(&y_bar[c & 1]);
            };
            if (producer && ob - 1 != P.op_first[2 * n]) wdone_dummy(ob - 1);                   // alignment dummy before the output ops
            if (tid == 0) issue_y(0);
            for (int c = 0; c < P.nC; ++c) {
                const int last_op = ob + min(2 * c + 1, P.nX - 1);
                const int f = 128 * c + frow;
                const bool valid = f < D;
                // this thread's 16 x values of the chunk (L2-resident since stage 0), issued before the accumulator wait
                float xv[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) xv[j] = (valid && jc0 + j < nvalid) ? __ldg(a.x + (row0 + jc0 + j) * D + f) : 0.f;
                const float bias = valid ? __ldg(a.theta + st.seg_b[0] + f) : 0.f;
                wait_y(c);
                if (tid == 0 && c + 1 < P.nC) issue_y(c + 1);
                stamp();
                if constexpr (SCORE) {
                    // loss_function (utils.py:346-375) terms of this thread's feature for its 16 jets; Q = clip_values(X_hat)
                    float y[16];
                    tmem_ld16(t_lane + T_Y0 + 64 * (c & 1) + jc0, y);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (valid && jc0 + j < nvalid) {
                            const float p = xv[j], qv = clipv(y[j] + bias), d = p - qv;
                            rec_acc[j] += d * d; sc_mae[j] += fabsf(d);
                            if (a.s_kld) sc_kl[j] += kld_term_fast(p, qv);
                            if (a.s_jsd) { const float mm = (p + qv) / 2.f; sc_js[j] += (kld_term_fast(p, mm) + kld_term_fast(qv, mm)) / 2.f; }
                                if (a.s_xhat) a.s_xhat[(row0 + jc0 + j) * D + f] = qv;
                        }
                    }
                } else
                // reconstruction loss (models.py:66,72-78) and its gradient through clip_values
                {
                    float dbs = 0.f;
                    float y[16];
                    tmem_ld16(t_lane + T_Y0 + 64 * (c & 1) + jc0, y);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float gy = 0.f;
                        if (valid) {
                            const float yr = y[j] + bias, d = clipv(yr) - xv[j], wj = s_w[jc0 + j];
                            rec_acc[j] += a.mse ? d * d : fabsf(d);
                            const float gg = a.mse ? wj * 2.f * d * inv_d : wj * (float)((d > 0.f) - (d < 0.f)) * inv_d;
                            gy = clip_pass(yr) ? gg : 0.f;
                        }
                        dbs += gy; y[j] = gy;
                    }
                    if (is_train) { store16(gy_hi, gy_lo, frow, jc0, y); db_at(2 + (c & 1), h, frow) = dbs; }
                }
                stamp();
                if (!is_train) {
                    tc_fence_before();
                    __syncthreads();                                    // accumulator reads done before chunk c+2 overwrites it
                    if (producer) wdone(last_op);
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
                if (producer) wdone(last_op);
                stamp();
                // dW_out^T chunk -> part (lanes = output features: coalesced rows of the Keras [in, D] kernel)
                for (int g = h; g < (st.kin >> 4); g += 4) {
                    float v[16];
                    tmem_ld16(t_lane + T_DW + 16 * g, v);
                    if (valid) put_dw(part + st.seg_w[0] + f, D, 16 * g, st.in, v);
                }
                if (valid && h == 0) put_db(part + st.seg_b[0] + f, 2 + (c & 1), frow);
                tc_fence_before();
                stamp();
            }
            const float rs = transpose_reduce16(rec_acc, lane);
            if (lane < 16) s_rec[q * 64 + jc0 + lane] = rs;
            if constexpr (SCORE) {                                       // s_db is free outside training: [3][4 quarters][64 jets]
                const float r1 = transpose_reduce16(sc_mae, lane), r2 = transpose_reduce16(sc_kl, lane), r3 = transpose_reduce16(sc_js, lane);
                if (lane < 16) { s_db[q * 64 + jc0 + lane] = r1; s_db[256 + q * 64 + jc0 + lane] = r2; s_db[512 + q * 64 + jc0 + lane] = r3; }
            }
            if (is_train) { epi_grad(so); publish(); }
            stamp();
        }
        // =========================================================== backward chain: stages 2n .. 1
        if (is_train) {
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
                float b_mu = 0.f, b_lv = 0.f;
                if (s == n + 1 && lane < L) { b_mu = __ldg(a.theta + P.st[n].seg_b[0] + lane); b_lv = __ldg(a.theta + P.st[n].seg_b[1] + lane); }
                wait_mma();
                if (producer) wdone(P.op_bwd[s]);
                epi_dw(s, s & 1);
                if (s != n + 1) {
                    epi_grad(s);
                } else if (q == 0) {
                    // ---- latent backward: g_z -> g_mu (rows 0..15), g_logvar (rows 16..31) of the latent tile
                    const float inv_l = 1.f / (float)L;
                    float dbm = 0.f, dbl = 0.f;
                    float gz[16], mu[16], lv[16];
                    tmem_ld16(t_lane + T_CH + jc0, gz);
                    tmem_ld16(t_lane + T_MU + jc0, mu);
                    tmem_ld16(t_lane + T_LV + jc0, lv);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float gm = 0.f, gl = 0.f;
                        if (lane < L) {
                            const float m = mu[j] + b_mu, v = lv[j] + b_lv, gk = a.beta * s_w[jc0 + j];
                            const float e_raw = expf(v), s_raw = expf(v * 0.5f);
                            gm = gz[j] + gk * m * inv_l;
                            gl = gk * (-(1.f - (clip_pass(e_raw) ? e_raw : 0.f)) * 0.5f * inv_l)
                               + gz[j] * s_eps[lane * 64 + jc0 + j] * (clip_pass(s_raw) ? s_raw * 0.5f : 0.f);
                        }
                        mu[j] = gm; lv[j] = gl; dbm += gm; dbl += gl;
                    }
                    if (lane < 16) {
                        store16(tile_hi(n), tile_lo(n), lane, jc0, mu);
                        store16(tile_hi(n), tile_lo(n), 16 + lane, jc0, lv);
                        db_at(n & 1, h, lane) = dbm; db_at(n & 1, h, 16 + lane) = dbl;
                    }
                }
                publish();
                stamp();
            }
            // ======================================================= stage 0: dW_0^T = G_0 x X, X re-staged from L2
            {
                const F2Stage& st = P.st[0];
                for (int c0 = 0; c0 < P.nX; c0 += 7) {
                    const int c1 = min(P.nX, c0 + 7);
                    float4 ra[2], rb[2];
                    ldx(c0, ra); ldx(c0 + 1 < c1 ? c0 + 1 : P.nX, rb);
                    auto chunk = [&](int c, float4 (&r)[2]) {
                        const int xs = (c - c0) % kF2XSlots;
                        if (c - c0 >= kF2XSlots) wait_x(xs);
                        stx(xs, r);
                        ldx(c + 2 < c1 ? c + 2 : P.nX, r);
                        publish();
                        if (tid == 0) {
                            tc_fence_after();
                            issue(tm + 64 * (c - c0), opnd_k(smem_u32(tile_hi(0)), P.tile_rows[0] * 128),
                                  opnd_mn(smem_u32(xslot(xs)), 8192, 8192), 4, 64, false);
                            mma_commit(&xbar[xs]);
                        }
                    };
                    for (int c = c0; c < c1; c += 2) {
                        chunk(c, ra);
                        if (c + 1 < c1) chunk(c + 1, rb);
                    }
                    for (int c = max(c0, c1 - kF2XSlots); c < c1; ++c) wait_x((c - c0) % kF2XSlots);
                    stamp();
                    if (32 * q < st.out) {
                        const bool valid = frow < st.out;
                        for (int g = h; g < (c1 - c0) * 4; g += 4) {
                            float v[16];
                            tmem_ld16(t_lane + 16 * g, v);
                            if (valid) put_dw(part + st.seg_w[0] + frow, st.out, 64 * c0 + 16 * g, D, v);
                        }
                        if (valid && h == 0 && c0 == 0) put_db(part + st.seg_b[0] + frow, 0, frow);
                    }
                    publish();
                }
            }
        }
        stamp();
        // =========================================================== per-jet losses (models.py:123-132) and batch sums
        if (tid == 0) { wget(nops - 1); gbase += nops; }
        __syncthreads();
        // trailing dummies / shrunk ops have no consumer; the producer frees their slots only after thread 0 has seen every
        // arrival of this tile (a slot refilled earlier could complete twice before thread 0 polls it: parity aliasing)
        if (producer) {
            for (int k = (is_train ? P.op_bwd[1] : n_fwd_ops - 1) + 1; k < nops; ++k) wdone_dummy(k);
            gbase += nops;
        }
        if (SCORE) {
            if (tid < nvalid) {
                const int j = tid;
                auto q4 = [&](const float* b) { return (b[j] + b[64 + j]) + (b[128 + j] + b[192 + j]); };
                if (a.s_mse) a.s_mse[row0 + j] = q4(s_rec) / (float)D;
                if (a.s_mae) a.s_mae[row0 + j] = q4(s_db) / (float)D;
                if (a.s_kld) a.s_kld[row0 + j] = q4(s_db + 256);
                if (a.s_jsd) a.s_jsd[row0 + j] = q4(s_db + 512);
                if (a.s_lat) a.s_lat[row0 + j] = isfinite(s_kld[j]) ? s_kld[j] : 0.f;       // latent_loss, utils.py:378-386
                if (a.s_fin) a.s_fin[row0 + j] = 1;                                         // clip_values leaves no non-finite X_hat
            }
        } else if (tid < 64) {
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
