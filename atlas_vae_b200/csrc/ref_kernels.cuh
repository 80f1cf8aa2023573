// fp32 CUDA-core kernels of the OE-VAE step: the reference-dtype mode (OEVAE_PREC_FP32) and the parity
// anchor for the tcgen05 paths.  Arithmetic follows /root/reference/OE-VAE/models.py (cited per kernel);
// nothing here is a translation of TensorFlow code -- the reference has no kernels of its own.
#pragma once
#include "common.cuh"

namespace oevae {

// ------------------------------------------------------------------------------------------------
// Generic 64x64x16 register-tiled SGEMM:  C[r][c] = sum_{t in split} Aop(r,t) * Bop(t,c)
//   Aop(r,t) = A[r*sAr + t*sAt],  Bop(t,c) = B[t*sBt + c*sBc]   (one stride of each is 1)
// A_TC / B_TC: the contraction index t is the contiguous one of that operand (decides the load mapping).
// Dense forward  (models.py:31-34,50-52): r=jet, t=in,  c=out   A=act  (t contiguous) B=W (c contiguous)
// dX             (tape, models.py:183)  : r=jet, t=out, c=in    A=g    (t contiguous) B=W (t contiguous)
// dW,db          (tape, models.py:183)  : r=in(+1 ones row), t=jet, c=out   A=act (r contiguous) B=g (c contiguous)
// ------------------------------------------------------------------------------------------------
struct GemmArgs {
    const float* A; const float* B; float* C;
    long sAr, sAt, sBt, sBc;
    int R, Cn, T;            // output rows, output cols, contraction length
    int ldc;
    // epilogue
    const float* bias;       // FWD: + bias[c]
    int relu;                // FWD: max(.,0)
    int accumulate;          // DX : += C
    const float* mask;       // DX : *= (mask[r*ldmask + c] > 0)
    int ldmask;
    int ones_row;            // DW : row r == R-1 of Aop is all ones (bias gradient; kernel,bias are contiguous)
    // contraction range [t_base, t_end); with chunk > 0 (DW) it is split over blockIdx.z = slot:
    // slot z covers group z / slots_per_group (group_rows rows each), chunk (z % slots_per_group) of `chunk`
    // rows, clipped to the group and to t_end; the output goes to C + z*slot_stride.
    int chunk, slots_per_group, group_rows;
    int t_base, t_end;
    long slot_stride;
};

template <bool A_TC, bool B_TC>
__global__ void __launch_bounds__(256) gemm64_kernel(GemmArgs g) {
    constexpr int BM = 64, BN = 64, BK = 16;
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int r0 = blockIdx.x * BM, c0 = blockIdx.y * BN;
    int t0 = g.t_base, t1 = g.t_end;
    float* C = g.C;
    if (g.chunk > 0) {
        int grp = (int)blockIdx.z / g.slots_per_group, ck = (int)blockIdx.z - grp * g.slots_per_group;
        t0 = grp * g.group_rows + ck * g.chunk;
        t1 = min(min(t0 + g.chunk, (grp + 1) * g.group_rows), g.t_end);
        C += (long)blockIdx.z * g.slot_stride;
    }
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const int lastA = g.ones_row ? g.R - 1 : -1;
    for (int tt = t0; tt < t1; tt += BK) {
        // ---- load A tile (BM x BK) into As[k][m]
        if (A_TC) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int m = (tid >> 4) + 16 * i, k = tid & 15;
                int r = r0 + m, t = tt + k;
                float v = 0.f;
                if (r < g.R && t < t1) v = (r == lastA) ? 1.f : __ldg(g.A + r * g.sAr + t * g.sAt);
                As[k][m] = v;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int m = tid & 63, k = (tid >> 6) + 4 * i;
                int r = r0 + m, t = tt + k;
                float v = 0.f;
                if (r < g.R && t < t1) v = (r == lastA) ? 1.f : __ldg(g.A + r * g.sAr + t * g.sAt);
                As[k][m] = v;
            }
        }
        // ---- load B tile (BK x BN) into Bs[k][n]
        if (B_TC) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int n = (tid >> 4) + 16 * i, k = tid & 15;
                int c = c0 + n, t = tt + k;
                float v = 0.f;
                if (c < g.Cn && t < t1) v = __ldg(g.B + t * g.sBt + c * g.sBc);
                Bs[k][n] = v;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int n = tid & 63, k = (tid >> 6) + 4 * i;
                int c = c0 + n, t = tt + k;
                float v = 0.f;
                if (c < g.Cn && t < t1) v = __ldg(g.B + t * g.sBt + c * g.sBc);
                Bs[k][n] = v;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int r = r0 + ty * 4 + i;
        if (r >= g.R) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int c = c0 + tx * 4 + j;
            if (c >= g.Cn) continue;
            float v = acc[i][j];
            if (g.bias) v += __ldg(g.bias + c);
            if (g.accumulate) v += C[(long)r * g.ldc + c];
            if (g.mask) v = (__ldg(g.mask + (long)r * g.ldmask + c) > 0.f) ? v : 0.f;
            if (g.relu) v = fmaxf(v, 0.f);
            C[(long)r * g.ldc + c] = v;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Sampling.call (models.py:9-15) + get_KLD_loss (models.py:81-85), one thread per jet row.
//   rows [0, n_rows) of mu/lv [n_rows, L];  z/eps_out written for rows < n_dec only
//   eps source: ext != nullptr -> ext[(eg*B + i)*L + l]  (eg = group if n_eps>1 else 0); else Philox
// ------------------------------------------------------------------------------------------------
struct LatentArgs {
    const float* mu; const float* lv; float* z; float* eps_out; float* kld;
    const float* eps_ext; int n_eps;
    long n_rows, n_dec, B; int L;
    unsigned long long seed, row_offset; unsigned int step_base; int per_group_step;
};

__global__ void latent_fwd_kernel(LatentArgs a) {
    long row = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= a.n_rows) return;
    const int L = a.L;
    const float* mu = a.mu + row * L;
    const float* lv = a.lv + row * L;
    float s = 0.f;
    for (int l = 0; l < L; ++l) {
        float m = mu[l], v = lv[l];
        float e = clipv(expf(v));
        s += 1.f + v - e - m * m;
    }
    a.kld[row] = -(s / (float)L) / 2.f;
    if (row >= a.n_dec) return;
    long grp = row / a.B, i = row - grp * a.B;
    float* z = a.z + row * L;
    float* eo = a.eps_out + row * L;
    for (int q = 0; q < (L + 3) / 4; ++q) {
        float n[4];
        if (a.eps_ext) {
            const float* e = a.eps_ext + ((a.n_eps > 1 ? grp : 0) * a.B + i) * L;
            for (int j = 0; j < 4; ++j) n[j] = (4 * q + j < L) ? e[4 * q + j] : 0.f;
        } else {
            philox_normal4(a.seed, a.row_offset + (unsigned long long)i, (unsigned)q,
                           a.step_base + (a.per_group_step ? (unsigned)grp : 0u), n);
        }
        for (int j = 0; j < 4 && 4 * q + j < L; ++j) {
            int l = 4 * q + j;
            float sg = clipv(expf(lv[l] * 0.5f));
            z[l] = mu[l] + sg * n[j];               // tf.random.normal(mean=mu, stddev=sigma) = eps*sigma + mu
            eo[l] = n[j];
        }
    }
}

// get_reconstruction_loss (models.py:72-78) on clip_values(reconstructed) (models.py:66): one warp per jet.
__global__ void rec_fwd_kernel(const float* __restrict__ x, const float* __restrict__ y_raw, float* __restrict__ rec,
                               long n_rows, int D, int mse) {
    long row = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (row >= n_rows) return;
    const float* xr = x + row * D; const float* yr = y_raw + row * D;
    float s = 0.f;
    for (int j = lane; j < D; j += 32) {
        float d = xr[j] - clipv(yr[j]);
        s += mse ? d * d : fabsf(d);
    }
    s = warp_sum(s);
    if (lane == 0) rec[row] = s / (float)D;
}

// get_losses / get_OE_loss combination (models.py:88-105,123-132) + the per-row coefficients of the backward
// pass (d total / d rec, d total / d kld of every pass).  Row groups: g0 = bkg full pass; then
//   OE 'KLD'      : g1 = OoD encoder-only
//   rec OE types  : n_groups==2: g1 = OoD full pass (second bkg pass shares eps with the first)
//                   n_groups==3: g1 = second bkg pass (fresh eps), g2 = OoD pass
struct CombineArgs {
    const float* rec; const float* kld; const float* w; const float* w_ood;
    float* g_rec; float* g_kld; float* per_jet; double* sums;
    long B; int n_groups; int oe_type; int have_ood;
    float beta, lamb, margin;
};

__global__ void combine_kernel(CombineArgs a) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    float lr_ = 0.f, lk = 0.f, lo = 0.f, tot = 0.f;
    if (i < a.B) {
        const long B = a.B;
        float w = a.w[i];
        float rec0 = a.rec[i], kld0 = a.kld[i];
        lr_ = rec0 * w;
        lk = kld0 * w * a.beta;
        float grec0 = w, gkld0 = w * a.beta;
        if (a.have_ood) {
            float wo = a.w_ood[i];
            float oe, coef;
            if (a.oe_type == 0) {
                float delta = kld0 - a.kld[B + i] + a.margin;
                oe = fmaxf(delta, 0.f);
                coef = a.lamb * wo * (delta > 0.f ? 1.f : 0.f);
                gkld0 += coef;
                if (a.g_kld) { a.g_kld[B + i] = -coef; a.g_rec[B + i] = 0.f; }
            } else {
                int go = a.n_groups - 1;
                float r2 = (a.n_groups == 3) ? a.rec[B + i] : rec0;
                float ro = a.rec[go * B + i];
                float dphi;
                if (a.oe_type == 1 || a.oe_type == 2) {
                    oe = 1.f / (1.f + expf(-(r2 - ro)));
                    dphi = oe * (1.f - oe);
                } else {
                    float delta = r2 - ro + a.margin;
                    oe = fmaxf(delta, 0.f);
                    dphi = delta > 0.f ? 1.f : 0.f;
                }
                coef = a.lamb * wo * dphi;
                if (a.g_rec) {
                    if (a.n_groups == 3) { a.g_rec[B + i] = coef; a.g_kld[B + i] = 0.f; }
                    else grec0 += coef;
                    a.g_rec[go * B + i] = -coef; a.g_kld[go * B + i] = 0.f;
                }
            }
            lo = oe * wo * a.lamb;
        }
        if (a.g_rec) { a.g_rec[i] = grec0; a.g_kld[i] = gkld0; }
        tot = lr_ + lk + lo;
        if (a.per_jet) {
            a.per_jet[i] = lr_; a.per_jet[B + i] = lk; a.per_jet[2 * B + i] = lo; a.per_jet[3 * B + i] = tot;
        }
    }
    if (a.sums) {                                       // keras.metrics.Mean numerators (models.py:187-190)
        __shared__ double sh[4][8];
        double v[4] = {lr_, lk, lo, tot};
        int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int k = 0; k < 4; ++k) { v[k] = warp_sum(v[k]); if (lane == 0) sh[k][wid] = v[k]; }
        __syncthreads();
        if (threadIdx.x < 4) {
            double s = 0;
            for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s += sh[threadIdx.x][q];
            atomicAdd(a.sums + threadIdx.x, s);
        }
    }
}

// backward of the reconstruction loss through clip_values: g_yraw overwrites y_raw in place.
__global__ void rec_bwd_kernel(const float* __restrict__ x, float* __restrict__ y_raw, const float* __restrict__ g_rec,
                               long n_rows, int D, int mse) {
    long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_rows * D) return;
    long row = idx / D;
    float yr = y_raw[idx];
    float y = clipv(yr), d = y - x[idx];
    float gr = g_rec[row];
    float gy = mse ? gr * 2.f * d / (float)D : gr * ((d > 0.f) - (d < 0.f)) / (float)D;
    y_raw[idx] = clip_pass(yr) ? gy : 0.f;
}

// backward of Sampling + KLD: g_mu, g_lv for every row; g_z only for rows < n_dec.
__global__ void latent_bwd_kernel(const float* __restrict__ mu, const float* __restrict__ lv, const float* __restrict__ eps,
                                  const float* __restrict__ g_z, const float* __restrict__ g_kld,
                                  float* __restrict__ g_mu, float* __restrict__ g_lv, long n_rows, long n_dec, int L) {
    long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_rows * L) return;
    long row = idx / L;
    float m = mu[idx], v = lv[idx], gk = g_kld[row];
    float e_raw = expf(v);
    float gm = gk * m / (float)L;
    float gl = gk * (-(1.f - (clip_pass(e_raw) ? e_raw : 0.f)) / (2.f * (float)L));
    if (row < n_dec) {
        float gz = g_z[idx];
        float s_raw = expf(v * 0.5f);
        gm += gz;
        gl += gz * eps[idx] * (clip_pass(s_raw) ? s_raw * 0.5f : 0.f);
    }
    g_mu[idx] = gm; g_lv[idx] = gl;
}

// sum of the split-K partial gradients (fixed order => deterministic), optional export, clip_values
// (models.py:185) and Keras-TF2.1 Adam (models.py:138,186; ResourceApplyAdam: eps outside the bias correction).
struct AdamArgs {
    const float* parts; int n_parts; long part_stride;     // parts==nullptr -> gradient comes from grad_in
    const float* grad_in; float* grad_out;                 // grad_out: write the raw summed gradient
    float* theta; float* m; float* v; long n;
    float alpha, b1, b2, eps; int apply;
};

__global__ void adam_kernel(AdamArgs a) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    float g;
    if (a.parts) {
        g = 0.f;
        for (int s = 0; s < a.n_parts; ++s) g += a.parts[(long)s * a.part_stride + i];
        if (a.grad_out) a.grad_out[i] = g;
    } else {
        g = a.grad_in[i];
    }
    if (!a.apply) return;
    g = clipv(g);
    float m = a.m[i], v = a.v[i];
    m += (g - m) * (1.f - a.b1);
    v += (g * g - v) * (1.f - a.b2);
    a.m[i] = m; a.v[i] = v;
    a.theta[i] -= (m * a.alpha) / (sqrtf(v) + a.eps);
}

// loss_function MSE/MAE/KLD/JSD (utils.py:346-375) between X_true and the n_iter-mean of the prediction
// (vae.py:166), plus the filtering flag (utils.py:532-538).  One warp per jet.
__device__ __forceinline__ float nan_to_num(float t) {
    if (isnan(t)) return 0.f;
    if (isinf(t)) return t > 0.f ? 3.402823466e+38f : -3.402823466e+38f;
    return t;
}
__device__ __forceinline__ float kld_term(float p, float q) { return nan_to_num(p * log2f(p / q)); }

struct MetricOut { float* mse; float* mae; float* mare; float* kld; float* jsd; float* xs; unsigned char* finite; float* xhat; };

// q = clip_in ? clip_values(ysum) : ysum * inv_iter   (ysum: raw reconstruction, n_iter sum, or a caller's Q)
__global__ void score_metrics_kernel(const float* __restrict__ x, const float* __restrict__ ysum, float inv_iter, int clip_in,
                                     long n_rows, int D, MetricOut o) {
    long row = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (row >= n_rows) return;
    const float* xr = x + row * D; const float* yr = ysum + row * D;
    float s2 = 0.f, s1 = 0.f, sr = 0.f, sk = 0.f, sj = 0.f, sx = 0.f; int ok = 1;
    for (int j = lane; j < D; j += 32) {
        float p = xr[j];
        float q = clip_in ? clipv(yr[j]) : yr[j] * inv_iter;
        if (o.xhat) o.xhat[row * D + j] = q;
        ok &= isfinite(q) ? 1 : 0;
        float d = p - q;
        s2 += d * d; s1 += fabsf(d);
        if (o.mare) sr += fabsf(d) / p;
        if (o.kld) sk += kld_term(p, q);
        if (o.jsd) { float mm = (p + q) / 2.f; sj += (kld_term(p, mm) + kld_term(q, mm)) / 2.f; }
        if (o.xs) sx += kld_term(p, p * q);
    }
    s2 = warp_sum(s2); s1 = warp_sum(s1); sr = warp_sum(sr); sk = warp_sum(sk); sj = warp_sum(sj); sx = warp_sum(sx);
    ok = __all_sync(0xffffffffu, ok);
    if (lane == 0) {
        if (o.mse) o.mse[row] = s2 / (float)D;
        if (o.mae) o.mae[row] = s1 / (float)D;
        if (o.mare) o.mare[row] = sr / (float)D;
        if (o.kld) o.kld[row] = sk;
        if (o.jsd) o.jsd[row] = sj;
        if (o.xs) o.xs[row] = sx;
        if (o.finite) o.finite[row] = (unsigned char)ok;
    }
}

// n_iter accumulation of clip_values(reconstructed): ysum (+)= clipv(y_raw)
__global__ void accum_clip_kernel(const float* __restrict__ y_raw, float* __restrict__ ysum, long n, int first) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float y = clipv(y_raw[i]);
    ysum[i] = first ? y : ysum[i] + y;
}

// latent_loss (utils.py:378-386): non-finite KLD -> 0
__global__ void latent_out_kernel(const float* __restrict__ kld, float* __restrict__ out, long n) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float k = kld[i];
    out[i] = isfinite(k) ? k : 0.f;
}

// apply_scaling for affine scalers (utils.py:406-417): sklearn's in-place float32 transform evaluated in fp64
__global__ void affine_kernel(const float* __restrict__ x, float* __restrict__ y, long N, int D,
                              const double* __restrict__ center, const double* __restrict__ scale) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * D) return;
    int c = (int)(i % D);
    double t = (double)x[i];
    if (center) t = (double)(float)(t - center[c]);
    y[i] = (float)(t / scale[c]);
}

}  // namespace oevae
