// Update kernel of the fused path: fixed-order sum of the per-CTA partial gradients, clip_values (models.py:185),
// Keras Adam (models.py:138,186) and the refresh of the pre-swizzled bf16 hi/lo weight image, one launch per step.
#pragma once
#include "umma.cuh"
#include "common.cuh"

namespace oevae {

// ------------------------------------------------------------------------------------------------
// Sum of the per-CTA partial gradients (fixed order), clip_values (models.py:185), Keras Adam (models.py:186) and
// refresh of the bf16 hi/lo weight images consumed by the fused kernels (fused2_train.cuh / fused3_train.cuh).
// ------------------------------------------------------------------------------------------------
// Early parameter groups.  The training kernel's gradient for the output layer is final when ~60 % of its tile has run, and the chain
// layers' when only the layer-0 products are left (fused3_train.cuh drain warps count themselves into ctr[0] / ctr[1]).  A block of the
// update kernel whose 256 parameters lie inside such a group does not wait for the training GRID to finish (griddepcontrol.wait) but for
// the group's counter to reach `target` (cumulative over the handle's launches, so nothing is ever reset): its reduce, its exchange with
// the peers (dp_kernels.cuh) and its Adam update run on the SMs the training kernel leaves idle, under the rest of the backward pass.
// Safe because the counter is complete only when EVERY CTA is past its last read of the group's weights / biases and its last write of the
// group's partial slice.  Blocks are mapped to parameter chunks in REVERSE (block 0 = last chunk): the blocks that become resident first
// are the output layer's.  ctr == nullptr: every block waits for the grid.
struct EarlyGroups {
    const unsigned int* ctr; unsigned int target; int dry;   // dry: poll, then wait for the grid all the same (A/B of the polling alone)
    long lo[2], hi[2];                                      // parameter ranges [lo, hi) of group 0 (output layer) and group 1 (chain layers)
    int fin;                                                // ctr[2] counts the CTAs that have done ALL their work (stores, reductions, loss sums):
                                                            // blocks outside groups 0 / 1 poll it instead of griddepcontrol.wait, which also waits
                                                            // for the grid's exit and memory flush (~1 us)
};
// returns true when the block need not wait for the grid: its chunk belongs to an early group and that group is complete, or every
// CTA of the training grid has reported itself finished; false: the caller waits for the grid (griddepcontrol.wait)
__device__ __forceinline__ bool early_wait(const EarlyGroups& e, long c0, long c1) {
    if (!e.ctr) return false;
    int g = -1;
    if (c0 >= e.lo[0] && c1 <= e.hi[0]) g = 0; else if (c0 >= e.lo[1] && c1 <= e.hi[1]) g = 1; else if (e.fin) g = 2;
    if (g < 0) return false;
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        unsigned int seen;
        for (;;) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(e.ctr + g) : "memory");
            if ((int)(seen - e.target) >= 0) break;
            __nanosleep(100);                               // ~290 resident pollers on one L2 sector: measured harmless down to 64 ns
            if (clock64() - t0 > 20000000000ll) __trap();   // ~10 s: the training kernel died; fail loudly instead of hanging
        }
    }
    __syncthreads();
    return !e.dry;
}

// Read-back of the step's running loss sums without a copy node in the stream: the block that owns chunk 0 (layer 0: it always waits
// for the whole training grid) stores the four doubles into page-locked host memory mapped into the device's address space.
struct SumsTrace { const double* src; double* dst_host; };
__device__ __forceinline__ void write_trace(const SumsTrace& t, int chunk) {
    if (t.dst_host && chunk == 0 && threadIdx.x < 4) t.dst_host[threadIdx.x] = __ldcg(t.src + threadIdx.x);
}

struct AdamImgArgs {
    EarlyGroups early; SumsTrace trace;
    const float* parts; int n_parts; long part_stride;
    const float* grad_in; float* grad_out;
    float* theta; float* m; float* v; long n;
    float alpha, b1, b2, eps; int apply;
    unsigned char* wimg; const int* pos;                    // pos: [4][n] byte offsets (fwd hi, fwd lo, dx hi, dx lo), -1 = none
};

// Fixed-order sum of the partial-gradient list for one block of kAdamCols parameters: thread (cx, sl) owns FOUR consecutive
// parameters (one 16-byte load per slice) and sums slices sl, sl + 8, ..., ten independent loads in flight (one L2 round trip
// for the 79 slices of a B = 5000 step); the eight slice sums meet in shared memory and are combined in a fixed tree by the
// thread that owns the parameter, so the result does not depend on scheduling.  One resident wave: ceil(P / 256) blocks of 512.
constexpr int kAdamCols = 256, kAdamSlices = 8, kAdamThreads = (kAdamCols / 4) * kAdamSlices;
__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
// returns (for threads tid < kAdamCols) the sum over the n_parts slices of parameter chunk * kAdamCols + tid
__device__ __forceinline__ float reduce_parts(const float* __restrict__ parts, int n_parts, long part_stride, float (*red)[kAdamCols], int chunk) {
    const int cx = threadIdx.x & (kAdamCols / 4 - 1), sl = threadIdx.x / (kAdamCols / 4);
    const long i4 = (long)chunk * kAdamCols + 4 * cx;            // rows are padded to a multiple of 4 floats (+ slack): always readable
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* src = parts + i4;
    for (int base = sl; base < n_parts; base += 10 * kAdamSlices) {
        float4 t[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) {
            const int sidx = base + kAdamSlices * k;
            t[k] = sidx < n_parts ? ldcg4(src + (long)sidx * part_stride) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        g.x += (((t[0].x + t[1].x) + (t[2].x + t[3].x)) + ((t[4].x + t[5].x) + (t[6].x + t[7].x))) + (t[8].x + t[9].x);
        g.y += (((t[0].y + t[1].y) + (t[2].y + t[3].y)) + ((t[4].y + t[5].y) + (t[6].y + t[7].y))) + (t[8].y + t[9].y);
        g.z += (((t[0].z + t[1].z) + (t[2].z + t[3].z)) + ((t[4].z + t[5].z) + (t[6].z + t[7].z))) + (t[8].z + t[9].z);
        g.w += (((t[0].w + t[1].w) + (t[2].w + t[3].w)) + ((t[4].w + t[5].w) + (t[6].w + t[7].w))) + (t[8].w + t[9].w);
    }
    *reinterpret_cast<float4*>(&red[sl][4 * cx]) = g;
    __syncthreads();
    const int t = threadIdx.x;
    return t < kAdamCols ? ((red[0][t] + red[1][t]) + (red[2][t] + red[3][t])) + ((red[4][t] + red[5][t]) + (red[6][t] + red[7][t])) : 0.f;
}

__global__ void __launch_bounds__(kAdamThreads, 2) adam_img_kernel(AdamImgArgs a) {
    __shared__ __align__(16) float red[kAdamSlices][kAdamCols];
    // programmatic dependent launch (no-ops otherwise): let the next training kernel begin its set-up, then wait for
    // the kernel that produced the partial gradients -- or, for an early parameter group, only for that group's counter
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int chunk = (int)gridDim.x - 1 - (int)blockIdx.x;
    const long c0 = (long)chunk * kAdamCols;
    if (!early_wait(a.early, c0, min(c0 + kAdamCols, a.n))) asm volatile("griddepcontrol.wait;" ::: "memory");
    const long i = c0 + threadIdx.x;
    const bool in = threadIdx.x < kAdamCols && i < a.n;
    float g = 0.f;
    if (a.parts) g = reduce_parts(a.parts, a.n_parts, a.part_stride, red, chunk);
    write_trace(a.trace, chunk);
    if (!in) return;
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
