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
