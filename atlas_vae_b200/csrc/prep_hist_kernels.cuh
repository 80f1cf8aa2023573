// K0 (constituent preprocessing) and K5 (cut + per-class histogram): HBM-bound byte/compare work.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>

namespace oevae {

// ------------------------------------------------------------------------------------------------
// K0: jets_sorting (utils.py:564-571) + keep n_const / zero-pad (utils.py:89-97) + n_dims slice
// (utils.py:115-119) + get_4v (utils.py:300-308) + affine scaler (utils.py:406-417), one warp per jet.
// Sort key: (fp32 bits of pT, constituent index): descending pT, ties in DESCENDING index order, i.e.
// np.argsort(pt, kind='stable')[::-1].  (The reference's default introsort leaves the order of exact non-zero
// ties platform-defined; ties inside the all-zero padding block do not change the output.)
// All arithmetic that feeds a comparison or the jet mass is done with explicit round-to-nearest
// intrinsics (no FMA contraction) and in NumPy's evaluation order so that pt/m are bit-identical.
// ------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

// Bitonic sort (descending) of 32 R keys held R per lane -- the network of the shared-memory loop in preprocess_kernel, unrolled.
// Element e = R lane + r (R consecutive keys per lane), so the steps with distance j < R exchange registers of one lane and only
// the steps with j >= R cross lanes (shuffle by j / R): 15 shuffle steps at every R, against 20 / 25 / 30 with e = 32 r + lane.
template <int R>
__device__ __forceinline__ void bitonic_regs_desc(unsigned long long (&x)[R], int lane) {
#pragma unroll
    for (int k = 2; k <= 32 * R; k <<= 1) {
        // direction of the element's k-block; for k < R it depends on r only (taken per register below)
        const bool desc_lane = ((lane * R) & k) == 0;
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < R) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int rp = r | j;
                    if (rp != r) {
                        const bool desc = k < R ? ((r & k) == 0) : desc_lane;
                        const unsigned long long a = x[r], b = x[rp];
                        const bool sw = desc ? (a < b) : (a > b);
                        x[r] = sw ? b : a; x[rp] = sw ? a : b;
                    }
                }
            } else {
                const bool keep_max = ((lane & (j / R)) == 0) == desc_lane;      // k > j >= R: the direction is the lane's
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, x[r], j / R);
                    x[r] = ((x[r] < o) == keep_max) ? o : x[r];          // equal keys are the all-zero padding: either is right
                }
            }
        }
    }
}
template <int R>
__device__ __forceinline__ void bitonic_smem_via_regs(unsigned long long* keys, int lane) {
    unsigned long long x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) x[r] = keys[R * lane + r];
    bitonic_regs_desc<R>(x, lane);
#pragma unroll
    for (int r = 0; r < R; ++r) keys[R * lane + r] = x[r];
    __syncwarp();
}

template <typename T>
__global__ void preprocess_kernel(const T* __restrict__ raw, long N, int n_raw, int n_pad, int n_const, int n_dims,
                                  const double* __restrict__ center, const double* __restrict__ scale, int do_sort,
                                  float* __restrict__ x_out, float* __restrict__ pt_out, float* __restrict__ m_out) {
    extern __shared__ unsigned char smem_raw[];
    const int warps = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // per warp: n_raw*4 floats + n_pad u64 keys
    const size_t per_warp = (size_t)n_pad * 8 + (size_t)n_raw * 16 + 16;      // + 16: the warps' regions start on different banks
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw + per_warp * wid);
    float* v = reinterpret_cast<float*>(keys + n_pad);
    long jet = (long)blockIdx.x * warps + wid;
    if (jet < N) {
    const T* src = raw + jet * 4L * n_raw;
    // 16-byte loads when the jet's row is 16-byte aligned (4 n_raw elements: always for fp32 from an aligned base, even n_raw for fp16)
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (4 * n_raw * (int)sizeof(T)) % 16 == 0) {
        constexpr int kPer = 16 / (int)sizeof(T);
        const int n16 = 4 * n_raw / kPer;
        const uint4* s16 = reinterpret_cast<const uint4*>(src);
        for (int i = lane; i < n16; i += 32) {
            const uint4 q = __ldg(s16 + i);
            if constexpr (sizeof(T) == 4) {
                *reinterpret_cast<uint4*>(v + 4 * i) = q;
            } else {
                const __half2* h = reinterpret_cast<const __half2*>(&q);
                const float2 a = __half22float2(h[0]), b = __half22float2(h[1]), c = __half22float2(h[2]), d = __half22float2(h[3]);
                *reinterpret_cast<float4*>(v + 8 * i) = make_float4(a.x, a.y, b.x, b.y);
                *reinterpret_cast<float4*>(v + 8 * i + 4) = make_float4(c.x, c.y, d.x, d.y);
            }
        }
    } else {
        for (int i = lane; i < 4 * n_raw; i += 32) v[i] = to_f32<T>(src[i]);
    }
    __syncwarp();
    // Sort key = (bits of pT) << 32 | (index + 1), descending: argsort(pT)[::-1] of utils.py:564-571 with ties (equal pT) resolved
    // towards the higher index.  Real jets fill 20-60 of the n_raw slots, the rest is zero padding, and every slot with pT == +0
    // has the key (0, index + 1): those keys are already ordered among themselves (descending index) and sort below every other
    // key.  So the slots with pT != 0 are compacted to the front (nv of them), the pT == 0 slots follow in descending index order,
    // and only the first nvp = 2^ceil(log2 nv) entries go through the bitonic network: the same permutation as the full sort at a
    // fraction of the passes (nv = 50: 21 passes over 64 keys instead of 36 over 256).
    auto key_of = [&](int c) -> unsigned long long {
        const float px = v[4 * c + 1], py = v[4 * c + 2];
        const float pt = __fsqrt_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)));
        return ((unsigned long long)__float_as_uint(pt) << 32) | (unsigned)(c + 1);
    };
    int nvp = 0;
    if (!do_sort) {
        // keep the stored order, e.g. get_4v on already sorted jets
        for (int c = lane; c < n_pad; c += 32) keys[c] = c < n_raw ? (unsigned long long)(c + 1) : 0ull;
    } else {
        const unsigned lt = (1u << lane) - 1u, gt = lane == 31 ? 0u : ~((2u << lane) - 1u);
        for (int c = n_raw + lane; c < n_pad; c += 32) keys[c] = 0ull;
        int nv = 0;
        unsigned zmask = 0;                                      // lane b keeps the pT == 0 mask of chunk b (n_raw <= 1024)
        for (int c0 = 0; c0 < n_raw; c0 += 32) {
            const int c = c0 + lane;
            const unsigned long long k = c < n_raw ? key_of(c) : 0ull;
            const bool nz = (k >> 32) != 0ull;
            const unsigned m = __ballot_sync(0xffffffffu, nz);
            const unsigned mz = __ballot_sync(0xffffffffu, c < n_raw && !nz);
            if (lane == (c0 >> 5)) zmask = mz;
            if (nz) keys[nv + __popc(m & lt)] = k;
            nv += __popc(m);
        }
        int zb = nv;
        for (int c0 = (n_raw - 1) & ~31; c0 >= 0; c0 -= 32) {
            const int c = c0 + lane;
            const unsigned m = n_raw <= 1024 ? __shfl_sync(0xffffffffu, zmask, c0 >> 5)
                                             : __ballot_sync(0xffffffffu, c < n_raw && (key_of(c) >> 32) == 0ull);
            const bool z = (m >> lane) & 1u;
            if (z) keys[zb + __popc(m & gt)] = (unsigned long long)(c + 1);
            zb += __popc(m);
        }
        if (nv > 1) { nvp = 32; while (nvp < nv) nvp <<= 1; }
    }
    __syncwarp();
    // bitonic sort of the first nvp keys, descending
    // (one compare-exchange per lane and step: t enumerates the nvp / 2 pairs (i, i | j) of a step, so no lane idles, and the
    // exchange is two selects instead of a divergent swap)
    // up to 256 keys the network runs in registers (no shared-memory round trips, no warp barriers)
    if (nvp == 32) { bitonic_smem_via_regs<1>(keys, lane); nvp = 0; }
    else if (nvp == 64) { bitonic_smem_via_regs<2>(keys, lane); nvp = 0; }
    else if (nvp == 128) { bitonic_smem_via_regs<4>(keys, lane); nvp = 0; }
    else if (nvp == 256) { bitonic_smem_via_regs<8>(keys, lane); nvp = 0; }
    for (int k = 2; k <= nvp; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = lane; t < (nvp >> 1); t += 32) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), p = i | j;
                const unsigned long long a = keys[i], b = keys[p];
                const bool sw = ((i & k) == 0) ? (a < b) : (a > b);
                keys[i] = sw ? b : a; keys[p] = sw ? a : b;
            }
            __syncwarp();
        }
    }
    const int D = n_dims * n_const, skip = 4 - n_dims;
    float* xo = x_out + jet * (long)D;
    for (int o = lane; o < D; o += 32) {
        int c, d;
        if (n_dims == 4) { c = o >> 2; d = o & 3; }                 // (a runtime division costs more than the gather itself)
        else { c = o / 3; d = o - c * 3 + skip; }
        float val = 0.f;
        if (c < n_raw) {
            unsigned idx = (unsigned)(keys[c] & 0xffffffffull);
            if (idx) val = v[4 * (idx - 1) + d];
        }
        if (scale) {
            double t = (double)val;
            if (center) t = (double)(float)(t - center[o]);
            val = (float)(t / scale[o]);
        }
        xo[o] = val;
    }
    }                                                            // jet < N
    if (pt_out || m_out) {
        // Sequential sum over the kept constituents in np.sum(axis=1) order: one lane per (jet of the block, 4-vector component), so
        // the block's first warp forms the sums of all its jets at once (a warp per jet kept 28 lanes idle for n_const steps).
        __syncthreads();
        if (wid == 0) {
            const int jl = lane >> 2, comp = lane & 3;
            const long jj = (long)blockIdx.x * warps + jl;
            const bool on = jl < warps && jj < N;
            float s = 0.f;
            if (on) {
                const unsigned long long* kj = reinterpret_cast<const unsigned long long*>(smem_raw + per_warp * jl);
                const float* vj = reinterpret_cast<const float*>(kj + n_pad);
                const int nk = min(n_const, n_raw);
#pragma unroll 4
                for (int c = 0; c < nk; ++c) {
                    const unsigned idx = (unsigned)(kj[c] & 0xffffffffull);
                    s = __fadd_rn(s, vj[4 * (idx - 1) + comp]);
                }
            }
            const int l0 = lane & ~3;
            const float E = __shfl_sync(0xffffffffu, s, l0), px = __shfl_sync(0xffffffffu, s, l0 + 1);
            const float py = __shfl_sync(0xffffffffu, s, l0 + 2), pz = __shfl_sync(0xffffffffu, s, l0 + 3);
            if (on && comp == 0) {
                const float px2 = __fmul_rn(px, px), py2 = __fmul_rn(py, py);
                if (pt_out) pt_out[jj] = __fsqrt_rn(__fadd_rn(px2, py2));
                if (m_out) {
                    const float m2 = __fsub_rn(__fsub_rn(__fsub_rn(__fmul_rn(E, E), px2), py2), __fmul_rn(pz, pz));
                    m_out[jj] = __fsqrt_rn(fmaxf(0.f, m2));
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// fp16 ingest: the reference's files hold the constituents as float16 (tools/merging.py:31,51-53), load_data widens them to float32
// (utils.py:89-97: jets_sorting returns np.float32) and apply_scaling transforms them (utils.py:406-417).  A host that keeps the
// sorted / truncated fp16 rows uploads half the bytes; this kernel is float32(raw) followed by the same fp64-evaluated affine map
// as preprocess_kernel / affine_kernel, eight elements per thread (one 16-byte load, two 16-byte stores).
// ------------------------------------------------------------------------------------------------
__global__ void ingest_f16_kernel(const __half* __restrict__ raw, float* __restrict__ x, long n_elems, int D,
                                  const double* __restrict__ center, const double* __restrict__ scale) {
    const long i8 = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i8 >= n_elems) return;
    float v[8];
    const bool vec = i8 + 8 <= n_elems && ((reinterpret_cast<uintptr_t>(raw + i8) & 15) == 0) && ((reinterpret_cast<uintptr_t>(x + i8) & 15) == 0);
    if (vec) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(raw + i8));
        const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float2 f = __half22float2(h[k]); v[2 * k] = f.x; v[2 * k + 1] = f.y; }
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = i8 + k < n_elems ? __half2float(raw[i8 + k]) : 0.f;
    }
    if (scale) {
        int c = (int)(i8 % D);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            double t = (double)v[k];
            if (center) t = (double)(float)(t - center[c]);
            v[k] = (float)(t / scale[c]);
            if (++c == D) c = 0;
        }
    }
    if (vec) {
        *reinterpret_cast<float4*>(x + i8) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(x + i8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) if (i8 + k < n_elems) x[i8 + k] = v[k];
    }
}

// ------------------------------------------------------------------------------------------------
// loss_mapping (plots.py:15-20): global min/max (decides the branch) and the elementwise map.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void atomic_min_double(double* addr, double val) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) <= val) break;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(val));
    } while (assumed != old);
}
__device__ __forceinline__ void atomic_max_double(double* addr, double val) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) >= val) break;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(val));
    } while (assumed != old);
}

// minmax[0] = min, minmax[1] = max, minmax[2] += number of NaN entries; caller initialises {+inf,-inf,0}
__global__ void minmax_kernel(const float* __restrict__ v, const unsigned char* __restrict__ keep, long N, double* minmax) {
    float lo = INFINITY, hi = -INFINITY; int nan = 0;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long)gridDim.x * blockDim.x) {
        if (keep && !keep[i]) continue;
        float x = v[i];
        if (isnan(x)) { nan++; continue; }
        lo = fminf(lo, x); hi = fmaxf(hi, x);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        nan += __shfl_xor_sync(0xffffffffu, nan, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (lo <= hi) { atomic_min_double(minmax, (double)lo); atomic_max_double(minmax + 1, (double)hi); }
        if (nan) atomicAdd(minmax + 2, (double)nan);
    }
}

__global__ void loss_map_kernel(float* __restrict__ v, long N, int branch) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float x = v[i], y;
    switch (branch) {
        case 0: y = x; break;
        case 1: y = x + 1.f; break;
        case 2: y = x / (fabsf(x) + 1.f); break;
        case 3: y = x / (fabsf(x) + 1.f) + 1.f; break;
        default: y = (x / (fabsf(x) + 1.f) + 1.f) / 2.f; break;
    }
    v[i] = y;
}

// ------------------------------------------------------------------------------------------------
// K5: make_cut mask (plots.py:259: X_loss > loss_cut) + per-class histogram of plot_distributions
// (plots.py:697-715; np.histogram semantics with float64 edges: bin k = [e_k, e_k+1), last bin closed,
// out-of-range and NaN values dropped).  Shared-memory privatised integer counts and fp64 weighted sums
// per CTA, flushed with one atomic per non-empty bin.
// Weight scalings of plots.py:706-712 (normalize / density): the per-jet weight becomes, in the reference's own rounding sequence,
//   w1 = float32(w * wscale)            `weights *= 100 / np.sum(...)`   (float32 array times float32 scalar)
//   w2 = float32(double(w1) / width_k)  `weights /= np.take(np.diff(bins), ...)`   (float32 array by float64 widths)
// and the bin value is the float64 sum of those float32 weights (pylab.hist sums them in float32 through a cumulative sum; the two
// agree to float32 rounding of the sum).
// ------------------------------------------------------------------------------------------------
__global__ void hist_kernel(const float* __restrict__ var, const float* __restrict__ w, const float* __restrict__ jzw,
                            const float* __restrict__ score, const unsigned char* __restrict__ keep, long N, float cut,
                            const double* __restrict__ edges, int n_edges, float wscale, int density,
                            unsigned long long* __restrict__ counts, double* __restrict__ wsums) {
    extern __shared__ unsigned char smem_raw[];
    const int nb = n_edges - 1;
    double* s_edges = reinterpret_cast<double*>(smem_raw);
    double* s_w = s_edges + n_edges;                       // [2*nb]
    unsigned int* s_c = reinterpret_cast<unsigned int*>(s_w + 2 * nb);   // [2*nb]
    for (int i = threadIdx.x; i < n_edges; i += blockDim.x) s_edges[i] = edges[i];
    for (int i = threadIdx.x; i < 2 * nb; i += blockDim.x) { s_w[i] = 0.0; s_c[i] = 0u; }
    __syncthreads();
    const double e0 = s_edges[0], e1 = s_edges[nb];
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long)gridDim.x * blockDim.x) {
        if (keep && !keep[i]) continue;
        if (score && !(score[i] > cut)) continue;          // NaN scores fail the cut
        float z = jzw[i];
        int cls = (z == -1.f) ? 0 : (z >= 0.f ? 1 : -1);
        if (cls < 0) continue;
        double x = (double)var[i];
        if (!(x >= e0 && x <= e1)) continue;               // also drops NaN
        int lo = 0, hi = nb;                               // largest k with edges[k] <= x  (searchsorted right - 1)
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (s_edges[mid] <= x) lo = mid; else hi = mid; }
        int bin = lo;                                      // x == last edge lands in the last bin (lo == nb-1)
        atomicAdd(&s_c[cls * nb + bin], 1u);
        if (w) {
            float wi = __fmul_rn(w[i], wscale);
            if (density) wi = (float)__ddiv_rn((double)wi, __dsub_rn(s_edges[bin + 1], s_edges[bin]));
            atomicAdd(&s_w[cls * nb + bin], (double)wi);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * nb; i += blockDim.x) {
        if (s_c[i]) {
            atomicAdd(&counts[i], (unsigned long long)s_c[i]);
            if (w) atomicAdd(&wsums[i], s_w[i]);
        }
    }
}

// fp64 sum of a float32 vector over the rows with keep[i] (nullable): np.sum(sample['weights']) of plots.py:707, accumulated ADDITIVELY
__global__ void wsum_kernel(const float* __restrict__ w, const unsigned char* __restrict__ keep, long N, double* __restrict__ out) {
    double s = 0.0;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long)gridDim.x * blockDim.x) if (!keep || keep[i]) s += (double)w[i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0 && s != 0.0) atomicAdd(out, s);
}

}  // namespace oevae
