// Weighted ROC curve and cut thresholds on the device (the step right after scoring):
//   get_rates       plots.py:228-232   sklearn.metrics.roc_curve(y_true, X_loss, pos_label=0, sample_weight=weights)
//   best_threshold  plots.py:235-247   argmax tpr/fpr over tpr > min_tpr
//   make_cut        plots.py:250-258   thresholds[argmin |fpr - bkg_eff|]
//   bump_scan grid  plots.py:272-282   searchsorted(fpr, eff_val, side='right')
// roc_curve's arithmetic (scikit-learn 1.9: _sort_inputs_and_compute_classification_thresholds, confusion_matrix_at_thresholds,
// roc_curve) is restated step by step so that fpr / tpr / thresholds are BIT-IDENTICAL to sklearn's:
//   1. samples with weight 0 are dropped (order-preserving compaction);
//   2. stable sort by DESCENDING score, ties in original order: an LSD radix sort (4 passes of 8 bits) on the bit-flipped,
//      order-preserving integer image of the float32 score, payload = sample index;
//   3. tps = cumsum(pos * w), fps = cumsum((1 - pos) * w) in float64 in np.cumsum's SEQUENTIAL order.  Floating-point addition
//      is not associative, so a parallel scan would round differently and drop_intermediate (step 5) would keep other points.
//      When every weight is an integer the partial sums are exact in any order and a parallel scan is used; otherwise one
//      thread walks the chain (two independent chains interleaved), fed through shared memory by the rest of its block;
//   4. one point per distinct score (last index of each run of ties);
//   5. drop_intermediate: keep k iff it is an end point or diff(fps, 2)[k-1] != 0 or diff(tps, 2)[k-1] != 0 (same float64
//      expression (f[k+1] - f[k]) - (f[k] - f[k-1]) as np.diff(f, 2));
//   6. prepend (0, 0, inf); fpr = fps / fps[-1], tpr = tps / tps[-1].
// HBM-bound integer / byte work: ~50 bytes per sample and radix pass; no tensor cores.
#pragma once
#include "common.cuh"

namespace oevae {
namespace roc {

constexpr int kTile = 4096, kThreads = 256, kItems = kTile / kThreads;       // one block = one tile of 4096 elements

// ---- order-preserving integer image of a float32 (ascending), -0.0 folded onto +0.0 like the float comparison
__device__ __forceinline__ uint32_t orderable(float f) {
    uint32_t u = __float_as_uint(f);
    if (u == 0x80000000u) u = 0u;
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// ================================================================================================ scans
// exclusive scan of uint32 [n] -> out [n], total in *total.  scan_tiles: per-tile sums; scan_block: in-place exclusive scan of the
// tile sums by one block; scan_apply: per-tile rescan + tile offset.
__global__ void __launch_bounds__(kThreads) scan_tiles_kernel(const uint32_t* __restrict__ in, long n, uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t red[kThreads / 32];
    const long base = (long)blockIdx.x * kTile;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) { const long i = base + k * kThreads + threadIdx.x; if (i < n) s += in[i]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < kThreads / 32; ++w) t += red[w]; tile_sums[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(1024) scan_block_kernel(uint32_t* __restrict__ data, long n, uint32_t* __restrict__ total) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long c0 = 0; c0 < n; c0 += 1024) {
        const long i = c0 + threadIdx.x;
        const uint32_t v = i < n ? data[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += t; }
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t y = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, y, o); if (lane >= o) y += t; }
            wsum[lane] = y;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t excl = carry + (warp ? wsum[warp - 1] : 0u) + x - v;
        if (i < n) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + wsum[31];
        __syncthreads();
    }
    if (threadIdx.x == 0 && total) *total = carry_s;
}
__global__ void __launch_bounds__(kThreads) scan_apply_kernel(const uint32_t* __restrict__ in, long n, const uint32_t* __restrict__ tile_offs,
                                                              uint32_t* __restrict__ out) {
    __shared__ uint32_t wsum[kThreads / 32];
    const long base = (long)blockIdx.x * kTile + (long)threadIdx.x * kItems;     // thread owns kItems CONSECUTIVE elements
    uint32_t v[kItems], s = 0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) { v[k] = base + k < n ? in[base + k] : 0u; s += v[k]; }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t x = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += t; }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    uint32_t off = tile_offs[blockIdx.x] + x - s;
    for (int w = 0; w < warp; ++w) off += wsum[w];
#pragma unroll
    for (int k = 0; k < kItems; ++k) { if (base + k < n) out[base + k] = off; off += v[k]; }
}

// ================================================================================================ radix sort (keys + index payload)
__global__ void __launch_bounds__(kThreads) rs_hist_kernel(const uint32_t* __restrict__ keys, long n, int shift, uint32_t* __restrict__ hist, int nblk) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const long base = (long)blockIdx.x * kTile;
#pragma unroll
    for (int k = 0; k < kItems; ++k) { const long i = base + k * kThreads + threadIdx.x; if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u); }
    __syncthreads();
    hist[(long)threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];              // digit-major: one exclusive scan gives the global offsets
}
// Stable scatter: warp w of the block owns elements [w * 512, w * 512 + 512) of the tile and walks them 32 at a time, so "earlier
// in the tile" = (earlier warp) or (same warp, earlier iteration) or (same iteration, lower lane).
__global__ void __launch_bounds__(kThreads) rs_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                              uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, long n, int shift,
                                                              const uint32_t* __restrict__ offs, int nblk) {
    constexpr int kWarps = kThreads / 32, kPerWarp = kTile / kWarps, kIter = kPerWarp / 32;
    __shared__ uint32_t cnt[kWarps][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kWarps * 256; i += kThreads) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const long base = (long)blockIdx.x * kTile + (long)warp * kPerWarp;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t key[kIter], pos[kIter];
#pragma unroll
    for (int j = 0; j < kIter; ++j) {
        const long i = base + j * 32 + lane;
        const bool on = i < n;
        key[j] = on ? keys_in[i] : 0u;
        const uint32_t d = on ? ((key[j] >> shift) & 255u) : (256u + lane);     // lanes past the end form singleton groups
        const uint32_t m = __match_any_sync(0xffffffffu, d);
        uint32_t before = 0;
        if (on) before = cnt[warp][d];
        __syncwarp();
        if (on && (m & lt) == 0u) cnt[warp][d] = before + __popc(m);            // the lowest lane of the group publishes the new count
        __syncwarp();
        pos[j] = before + __popc(m & lt);
    }
    __syncthreads();
    {   // digit d (one per thread): global offset of this tile's first element with that digit, then the warps' sub-ranges
        const int d = threadIdx.x;
        uint32_t run = offs[(long)d * nblk + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { const uint32_t t = cnt[w][d]; cnt[w][d] = run; run += t; }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kIter; ++j) {
        const long i = base + j * 32 + lane;
        if (i < n) {
            const uint32_t dst = cnt[warp][(key[j] >> shift) & 255u] + pos[j];
            keys_out[dst] = key[j];
            vals_out[dst] = vals_in[i];
        }
    }
}

__global__ void u8_flag_kernel(const unsigned char* __restrict__ f, long n, uint32_t* __restrict__ flags) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = f[i] ? 1u : 0u;
}
// flagged samples, compacted in order: keys = ascending image of the score, vals = sample index
__global__ void compact_score_keys_kernel(const float* __restrict__ score, long n, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos,
                                          uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flags[i]) { keys[pos[i]] = orderable(score[i]); vals[pos[i]] = (uint32_t)i; }
}
__global__ void gather_u32_kernel(const uint32_t* __restrict__ idx, const uint32_t* __restrict__ src, long n, uint32_t* __restrict__ dst) {
    const long p = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) dst[p] = src[idx[p]];
}
// ascending keys for a plain argsort (payload = index), and the uint32 -> int64 widening of the sorted payload
__global__ void asc_keys_kernel(const float* __restrict__ v, long n, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { keys[i] = orderable(v[i]); vals[i] = (uint32_t)i; }
}
__global__ void widen_idx_kernel(const uint32_t* __restrict__ in, long n, long long* __restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (long long)in[i];
}

// ================================================================================================ ROC stages
// flags[i] = weight != 0 (sklearn drops zero-weight samples before sorting)
__global__ void nonzero_flag_kernel(const float* __restrict__ w, long n, uint32_t* __restrict__ flags) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = w[i] != 0.f ? 1u : 0u;
}
// keys[p] = ~orderable(score[i]) (ascending radix order = descending score), vals[p] = i for the kept samples; bad |= non-finite score
__global__ void build_keys_kernel(const float* __restrict__ score, long n, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos,
                                  uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, int* __restrict__ bad) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flags && !flags[i]) return;
    const float s = score[i];
    if (!isfinite(s)) atomicOr(bad, 1);
    const long p = flags ? (long)pos[i] : i;
    keys[p] = ~orderable(s);
    vals[p] = (uint32_t)i;
}
// sorted position p: a[p] = pos ? w : 0, b[p] = pos ? 0 : w (float64, exact widening), s[p] = score; bad |= 2 when a weight is not an integer
__global__ void gather_sorted_kernel(const uint32_t* __restrict__ vals, long m, const float* __restrict__ score, const unsigned char* __restrict__ y_true,
                                     int pos_label, const float* __restrict__ w, double* __restrict__ a, double* __restrict__ b, float* __restrict__ s,
                                     int* __restrict__ bad) {
    const long p = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= m) return;
    const uint32_t i = vals[p];
    const float wi = w ? w[i] : 1.f;
    if (wi != floorf(wi) || fabsf(wi) > 16777216.f) atomicOr(bad, 2);
    const bool is_pos = (int)y_true[i] == pos_label;
    a[p] = is_pos ? (double)wi : 0.0;
    b[p] = is_pos ? 0.0 : (double)wi;
    s[p] = score[i];
}
// np.cumsum order: one thread adds, the block streams tiles through shared memory (double-buffered)
__global__ void __launch_bounds__(kThreads) seq_cumsum2_kernel(double* __restrict__ a, double* __restrict__ b, long n) {
    constexpr int T = 1024;                                                  // 2 x 2 x 1024 doubles = 32 KB of static shared memory
    __shared__ double sa[2][T], sb[2][T];
    double ra = 0.0, rb = 0.0;
    const long ntile = (n + T - 1) / T;
    auto load = [&](long t, int buf) {
        for (int k = threadIdx.x; k < T; k += kThreads) { const long i = t * T + k; sa[buf][k] = i < n ? a[i] : 0.0; sb[buf][k] = i < n ? b[i] : 0.0; }
    };
    load(0, 0);
    __syncthreads();
    for (long t = 0; t < ntile; ++t) {
        const int buf = (int)(t & 1);
        if (threadIdx.x == 0) {
#pragma unroll 8
            for (int k = 0; k < T; ++k) { ra += sa[buf][k]; sa[buf][k] = ra; rb += sb[buf][k]; sb[buf][k] = rb; }
        } else if (threadIdx.x >= 32 && t + 1 < ntile) {
            for (int k = threadIdx.x - 32; k < T; k += kThreads - 32) { const long i = (t + 1) * T + k; sa[buf ^ 1][k] = i < n ? a[i] : 0.0; sb[buf ^ 1][k] = i < n ? b[i] : 0.0; }
        }
        __syncthreads();
        for (int k = threadIdx.x; k < T; k += kThreads) { const long i = t * T + k; if (i < n) { a[i] = sa[buf][k]; b[i] = sb[buf][k]; } }
        __syncthreads();
    }
}
// exact (integer-valued) float64 inclusive scan, three kernels like the uint32 scan
__global__ void __launch_bounds__(kThreads) dscan_tiles_kernel(const double* __restrict__ a, const double* __restrict__ b, long n, double* __restrict__ ta, double* __restrict__ tb) {
    __shared__ double ra[kThreads / 32], rb[kThreads / 32];
    const long base = (long)blockIdx.x * kTile;
    double x = 0.0, y = 0.0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) { const long i = base + k * kThreads + threadIdx.x; if (i < n) { x += a[i]; y += b[i]; } }
    x = warp_sum(x); y = warp_sum(y);
    if ((threadIdx.x & 31) == 0) { ra[threadIdx.x >> 5] = x; rb[threadIdx.x >> 5] = y; }
    __syncthreads();
    if (threadIdx.x == 0) { double s = 0.0, t = 0.0; for (int w = 0; w < kThreads / 32; ++w) { s += ra[w]; t += rb[w]; } ta[blockIdx.x] = s; tb[blockIdx.x] = t; }
}
__global__ void dscan_block_kernel(double* __restrict__ ta, double* __restrict__ tb, long n) {      // exclusive, one thread (n = tiles: a few thousand)
    if (threadIdx.x == 0 && blockIdx.x == 0) { double s = 0.0, t = 0.0; for (long i = 0; i < n; ++i) { const double x = ta[i], y = tb[i]; ta[i] = s; tb[i] = t; s += x; t += y; } }
}
__global__ void __launch_bounds__(kThreads) dscan_apply_kernel(double* __restrict__ a, double* __restrict__ b, long n, const double* __restrict__ ta, const double* __restrict__ tb) {
    __shared__ double wa[kThreads / 32], wb[kThreads / 32];
    const long base = (long)blockIdx.x * kTile + (long)threadIdx.x * kItems;
    double va[kItems], vb[kItems], sa = 0.0, sb = 0.0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) { va[k] = base + k < n ? a[base + k] : 0.0; vb[k] = base + k < n ? b[base + k] : 0.0; sa += va[k]; sb += vb[k]; }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double xa = sa, xb = sb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(0xffffffffu, xa, o), u = __shfl_up_sync(0xffffffffu, xb, o); if (lane >= o) { xa += t; xb += u; } }
    if (lane == 31) { wa[warp] = xa; wb[warp] = xb; }
    __syncthreads();
    double oa = ta[blockIdx.x] + xa - sa, ob = tb[blockIdx.x] + xb - sb;
    for (int w = 0; w < warp; ++w) { oa += wa[w]; ob += wb[w]; }
#pragma unroll
    for (int k = 0; k < kItems; ++k) { oa += va[k]; ob += vb[k]; if (base + k < n) { a[base + k] = oa; b[base + k] = ob; } }
}
// one point per distinct score: flag the last index of every run of equal scores
__global__ void distinct_flag_kernel(const float* __restrict__ s, long m, uint32_t* __restrict__ flags) {
    const long p = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < m) flags[p] = (p == m - 1 || s[p] != s[p + 1]) ? 1u : 0u;
}
__global__ void take_points_kernel(const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos, long m, const double* __restrict__ ca,
                                   const double* __restrict__ cb, const float* __restrict__ s, double* __restrict__ tps, double* __restrict__ fps, double* __restrict__ thr) {
    const long p = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < m && flags[p]) { const long k = pos[p]; tps[k] = ca[p]; fps[k] = cb[p]; thr[k] = (double)s[p]; }
}
// drop_intermediate: keep[k] = end point or a corner in fps or in tps (np.diff(., 2) != 0)
__global__ void corner_flag_kernel(const double* __restrict__ fps, const double* __restrict__ tps, long m, uint32_t* __restrict__ keep) {
    const long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    bool kp = k == 0 || k == m - 1;
    if (!kp) {
        const double df = __dsub_rn(__dsub_rn(fps[k + 1], fps[k]), __dsub_rn(fps[k], fps[k - 1]));
        const double dt = __dsub_rn(__dsub_rn(tps[k + 1], tps[k]), __dsub_rn(tps[k], tps[k - 1]));
        kp = df != 0.0 || dt != 0.0;
    }
    keep[k] = kp ? 1u : 0u;
}
// out[0] = (0, 0, inf); out[1 + j] = (fps / fps_last, tps / tps_last, thr) of the kept points
__global__ void finish_kernel(const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos, long m, const double* __restrict__ fps,
                              const double* __restrict__ tps, const double* __restrict__ thr, double* __restrict__ fpr_out, double* __restrict__ tpr_out,
                              double* __restrict__ thr_out) {
    const long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const double fl = fps[m - 1], tl = tps[m - 1];
    if (k == 0) { fpr_out[0] = fl > 0.0 ? 0.0 : NAN; tpr_out[0] = tl > 0.0 ? 0.0 : NAN; thr_out[0] = INFINITY; }   // a class without samples: all-NaN rate (sklearn warns)
    if (k >= m || (keep && !keep[k])) return;
    const long j = 1 + (keep ? (long)pos[k] : k);
    fpr_out[j] = fl > 0.0 ? __ddiv_rn(fps[k], fl) : NAN;
    tpr_out[j] = tl > 0.0 ? __ddiv_rn(tps[k], tl) : NAN;
    thr_out[j] = thr[k];
}

// ================================================================================================ cut thresholds from the curve
// get_rates (plots.py:230): points with fpr != 0, rates in percent.  Block e < n_eff: argmin |100 fpr - eff[e]| (make_cut, plots.py:256-258);
// block n_eff: argmax 100 tpr / (100 fpr) over 100 tpr > min_tpr (best_threshold 'gain', plots.py:237-246).  First occurrence wins, as
// np.argmin / np.argmax.  out[e] = {threshold, value (|fpr - eff| or the gain), index in the curve (as a double), 100 fpr there}
__global__ void __launch_bounds__(1024) roc_cuts_kernel(const double* __restrict__ fpr, const double* __restrict__ tpr, const double* __restrict__ thr, long n,
                                                        const double* __restrict__ eff, int n_eff, double min_tpr, double* __restrict__ out) {
    __shared__ double bv[32]; __shared__ long bi[32];
    const int e = blockIdx.x;
    const bool gain = e == n_eff;
    const double target = gain ? 0.0 : eff[e];
    double best = gain ? -INFINITY : INFINITY; long besti = -1;
    for (long i = threadIdx.x; i < n; i += blockDim.x) {
        const double f = fpr[i];
        if (!(f != 0.0)) continue;                                        // fpr != 0 (NaN != 0 is true in NumPy; a NaN curve has no cut)
        const double F = __dmul_rn(100.0, f), T = __dmul_rn(100.0, tpr[i]);
        // a thread visits its indices in increasing order, so a strict comparison keeps the first occurrence
        if (gain) {
            if (!(T > min_tpr)) continue;
            const double v = __ddiv_rn(T, F);
            if (besti < 0 || v > best) { best = v; besti = i; }
        } else {
            const double v = fabs(__dsub_rn(F, target));
            if (besti < 0 || v < best) { best = v; besti = i; }
        }
    }
    auto better = [&](double v, long i, double w, long j) {              // (v, i) beats (w, j)?
        if (i < 0) return false;
        if (j < 0) return true;
        if (gain ? v > w : v < w) return true;
        return v == w && i < j;
    };
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o); const long oi = __shfl_xor_sync(0xffffffffu, besti, o);
        if (better(ov, oi, best, besti)) { best = ov; besti = oi; }
    }
    if ((threadIdx.x & 31) == 0) { bv[threadIdx.x >> 5] = best; bi[threadIdx.x >> 5] = besti; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) if (better(bv[w], bi[w], best, besti)) { best = bv[w]; besti = bi[w]; }
        out[4 * e + 0] = besti >= 0 ? thr[besti] : NAN;
        out[4 * e + 1] = best;
        out[4 * e + 2] = (double)besti;                                   // index in the unfiltered curve (-1: no point qualifies)
        out[4 * e + 3] = besti >= 0 ? __dmul_rn(100.0, fpr[besti]) : NAN;
    }
}
// np.minimum(np.searchsorted(eff, v, side='right'), len(eff) - 1) over the fpr != 0 points (bump_scan, plots.py:282): eff = 100 fpr is
// non-decreasing; returns the threshold at that index.  One thread per query.
__global__ void roc_searchsorted_kernel(const double* __restrict__ fpr, const double* __restrict__ thr, long n, const double* __restrict__ q, int nq,
                                        double* __restrict__ thr_out, double* __restrict__ eff_out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq) return;
    const double v = q[t];
    long lo = 0, hi = n;                                                  // first: number of leading points with fpr == 0 (get_rates drops them)
    while (lo < hi) { const long mid = (lo + hi) >> 1; if (fpr[mid] > 0.0) hi = mid; else lo = mid + 1; }
    hi = n;                                                               // first index in [first, n) with 100 fpr > v
    while (lo < hi) { const long mid = (lo + hi) >> 1; if (__dmul_rn(100.0, fpr[mid]) > v) hi = mid; else lo = mid + 1; }
    const long idx = lo < n ? lo : n - 1;
    thr_out[t] = thr[idx];
    eff_out[t] = __dmul_rn(100.0, fpr[idx]);
}

}  // namespace roc

// ================================================================================================ OoD <-> background pairing
// OoD_pairing (OE-VAE/utils.py:193-227): every background jet is paired with ONE out-of-distribution jet drawn uniformly from the OoD
// jets inside an (m, pT) window around it -- m in [m - wm/2, m + wm/2) on the m-SORTED OoD sample (two np.searchsorted, side 'left'),
// pT in [pT - wp/2, pT + wp/2] -- starting at 10 x 10 GeV and doubling the pT width, then the m width, alternately, until the window
// is not empty (utils.py:203-210, same order).  One warp per background jet: the lanes scan the m-window of the sorted pT array
// (coalesced), count the matches, then fetch the k-th one with k = floor(u * count), u from Philox4x32-10 keyed by (seed, jet).
// The reference draws with np.random.choice under fork(), which cannot be reproduced; the candidate SET is identical (tested against
// the reference's own function body) and the draw is uniform over it.
namespace pair {

__device__ __forceinline__ long lower_bound_f(const float* __restrict__ a, long n, double v) {   // np.searchsorted(a, v) (side 'left'), float32 a vs float64 v
    long lo = 0, hi = n;
    while (lo < hi) { const long mid = (lo + hi) >> 1; if ((double)a[mid] < v) lo = mid + 1; else hi = mid; }
    return lo;
}

__global__ void __launch_bounds__(256) ood_pairing_kernel(const float* __restrict__ m_bkg, const float* __restrict__ pt_bkg, long n_bkg,
                                                          const float* __restrict__ m_ood, const float* __restrict__ pt_ood, long n_ood,
                                                          unsigned long long seed, int pick_first, long long* __restrict__ idx_out,
                                                          float* __restrict__ widths_out /* nullable [n_bkg][2] */) {
    const long jet = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (jet >= n_bkg) return;
    // the widths are Python floats in the reference (10 -> 20.0 -> ...): float64 arithmetic on float32 jet values
    const double mv = (double)m_bkg[jet], pv = (double)pt_bkg[jet];
    double wm = 10.0, wp = 10.0;
    long i0 = 0, i1 = 0; int cnt = 0;
    for (int it = 0; it < 4096; ++it) {                                  // widths double: 4096 rounds never happen on finite inputs
        i0 = lower_bound_f(m_ood, n_ood, mv - wm / 2); i1 = lower_bound_f(m_ood, n_ood, mv + wm / 2);
        const double plo = pv - wp / 2, phi = pv + wp / 2;
        int c = 0;
        for (long i = i0 + lane; i < i1; i += 32) { const double p = (double)pt_ood[i]; c += (p >= plo && p <= phi) ? 1 : 0; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        cnt = c;
        if (cnt > 0) break;
        if ((it & 1) == 0) wp *= 2; else wm *= 2;                        // pT first, then m (utils.py:206,209)
    }
    if (cnt == 0) { if (lane == 0) idx_out[jet] = -1; return; }           // empty OoD sample (the reference would loop forever)
    // k-th match, k uniform in [0, cnt)
    u32x4 r = philox4x32_10(u32x4{(uint32_t)jet, (uint32_t)((unsigned long long)jet >> 32), 0x0f1u, 0u}, (uint32_t)seed, (uint32_t)(seed >> 32));
    int k = pick_first ? 0 : (int)(((unsigned long long)r.x * (unsigned long long)cnt) >> 32);
    const double plo = pv - wp / 2, phi = pv + wp / 2;
    long found = -1;
    for (long base = i0; base < i1 && found < 0; base += 32) {
        const long i = base + lane;
        const bool hit = i < i1 && (double)pt_ood[i] >= plo && (double)pt_ood[i] <= phi;
        const unsigned mask = __ballot_sync(0xffffffffu, hit);
        const int c = __popc(mask);
        if (k < c) {                                                      // the k-th set bit of mask
            unsigned mm = mask;
            for (int t = 0; t < k; ++t) mm &= mm - 1;
            found = base + (__ffs(mm) - 1);
        } else k -= c;
    }
    if (lane == 0) {
        idx_out[jet] = found;
        if (widths_out) { widths_out[2 * jet] = (float)wm; widths_out[2 * jet + 1] = (float)wp; }
    }
}

}  // namespace pair

// ================================================================================================ mass decorrelation
// mass_deco (OE-VAE/plots.py:54-85): the score of every jet is replaced by its empirical CDF value among the BACKGROUND jets of the same
// (mass, pT) cell, which flattens the score in every cell.  The bin construction (get_bins, plots.py:54-63: 100 log-spaced edges,
// neighbouring bins merged until each holds >= 2 jets) is host logic on <= 100 counts; the device supplies the counts
// (digitize_count_kernel), the cell of every jet (cell_index_kernel), the background scores sorted by (cell, score) (the radix sort
// above, twice: LSD over the composite key) and the CDF evaluation (cdf_apply_kernel), which restates cum_distribution (plots.py:64-66:
// np.unique + cumsum / len, a 0 prepended to values and counts) and scipy.interpolate.interp1d's linear rule
//   y = f32((x - x_lo) / (x_hi - x_lo)) * y_hi + f32((x_hi - x) / (x_hi - x_lo)) * y_lo   (SciPy 1.18),   0 below the first value, 1 above the last
// without materialising the per-cell unique arrays: inside the cell's sorted segment, pos = first score >= x gives x_hi, the end of
// its run of ties gives y_hi = (#scores <= x_hi) / len, and pos itself gives x_lo (the previous score, or the prepended 0) and y_lo.
namespace deco {

// np.clip(np.digitize(x, edges), 1, len(edges) - 1) - 1   (digitize = searchsorted(edges, x, 'right'), float64 compare)
__device__ __forceinline__ int clip_digitize(double x, const double* __restrict__ e, int ne) {
    int lo = 0, hi = ne;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (e[mid] <= x) lo = mid + 1; else hi = mid; }
    return min(max(lo, 1), ne - 1) - 1;
}
__global__ void digitize_count_kernel(const float* __restrict__ var, const unsigned char* __restrict__ keep, long n, const double* __restrict__ edges, int ne,
                                      unsigned int* __restrict__ counts) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        if (!keep || keep[i]) atomicAdd(&counts[clip_digitize((double)var[i], edges, ne)], 1u);
}
// cell = m_bin * 256 + pt_bin, the pT edges being those of the jet's mass bin (pt_off[m_bin] .. pt_off[m_bin + 1] in pt_edges)
__global__ void cell_index_kernel(const float* __restrict__ m, const float* __restrict__ pt, long n, const double* __restrict__ m_edges, int n_me,
                                  const double* __restrict__ pt_edges, const int* __restrict__ pt_off, uint32_t* __restrict__ cell) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int mi = clip_digitize((double)m[i], m_edges, n_me);
    const int o = pt_off[mi], npe = pt_off[mi + 1] - o;
    cell[i] = (uint32_t)(mi * 256 + clip_digitize((double)pt[i], pt_edges + o, npe));
}
__global__ void cell_count_kernel(const uint32_t* __restrict__ cell, long n, uint32_t* __restrict__ counts) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicAdd(&counts[cell[i]], 1u);
}
// keys for the second (cell) pass of the composite sort: keys[p] = cell[order[p]], vals[p] = order[p]
__global__ void gather_cell_keys_kernel(const long long* __restrict__ order, const uint32_t* __restrict__ cell, long n, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const long p = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) { const long long i = order[p]; keys[p] = cell[i]; vals[p] = (uint32_t)i; }
}
__global__ void gather_f32_kernel(const uint32_t* __restrict__ idx, const float* __restrict__ src, long n, float* __restrict__ dst) {
    const long p = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) dst[p] = src[idx[p]];
}
// loss[i] <- CDF of its cell at loss[i]; sorted: background scores ordered by (cell, score); start[c] .. start[c + 1]: cell c's segment
__global__ void cdf_apply_kernel(float* __restrict__ loss, const uint32_t* __restrict__ cell, long n, const float* __restrict__ sorted, const uint32_t* __restrict__ start) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = cell[i];
    const long s0 = start[c], s1 = start[c + 1], len = s1 - s0;
    const float x = loss[i];
    if (len <= 0) { loss[i] = NAN; return; }                              // no background jet in the cell (the reference raises)
    const float last = sorted[s1 - 1];
    if (x < 0.f) { loss[i] = 0.f; return; }                               // below values[0] = 0 (the prepended point)
    if (x > fmaxf(last, 0.f)) { loss[i] = 1.f; return; }                  // above the last value
    long lo = s0, hi = s1;                                                // pos = first index with sorted[pos] >= x
    while (lo < hi) { const long mid = (lo + hi) >> 1; if (sorted[mid] < x) lo = mid + 1; else hi = mid; }
    // (x > last was excluded, so pos < s1 unless every score is negative, which loss_mapping rules out)
    const long pos = lo < s1 ? lo : s1 - 1;
    const float x_hi = sorted[pos];
    long l2 = pos, h2 = s1;                                               // end of the run of ties of x_hi
    while (l2 < h2) { const long mid = (l2 + h2) >> 1; if (sorted[mid] <= x_hi) l2 = mid + 1; else h2 = mid; }
    const float x_lo = pos > s0 ? sorted[pos - 1] : 0.f;
    const double y_lo = __ddiv_rn((double)(pos - s0), (double)len), y_hi = __ddiv_rn((double)(l2 - s0), (double)len);
    // SciPy 1.18 _call_linear: ((x - x_lo) / (x_hi - x_lo)) * y_hi + ((x_hi - x) / (x_hi - x_lo)) * y_lo, the two ratios in float32
    const float dx = __fsub_rn(x_hi, x_lo);
    const float a = __fdiv_rn(__fsub_rn(x, x_lo), dx), b = __fdiv_rn(__fsub_rn(x_hi, x), dx);
    loss[i] = (float)__dadd_rn(__dmul_rn((double)a, y_hi), __dmul_rn((double)b, y_lo));
}

}  // namespace deco
}  // namespace oevae
