// Shared device helpers for the OE-VAE kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

namespace oevae {

constexpr float kClip = 1e6f;               // clip_values max_val (models.py:12,66,84,185)

// clip_values (models.py:243-246): non-finite -> 0, then clamp to [-M, M]
__device__ __forceinline__ float clipv(float t) {
    t = isfinite(t) ? t : 0.f;
    return fminf(fmaxf(t, -kClip), kClip);
}
// gradient mask of clip_values: tf.where routes to the finite branch, clip_by_value passes on [-M, M]
__device__ __forceinline__ bool clip_pass(float t) { return isfinite(t) && fabsf(t) <= kClip; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- Philox4x32-10 (Salmon et al. SC'11), the noise spec of oracle/philox.py -------------------
struct u32x4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ u32x4 philox4x32_10(u32x4 c, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c.x, p1 = (uint64_t)M1 * c.z;
        u32x4 n;
        n.x = (uint32_t)(p1 >> 32) ^ c.y ^ k0;
        n.y = (uint32_t)p1;
        n.z = (uint32_t)(p0 >> 32) ^ c.w ^ k1;
        n.w = (uint32_t)p0;
        c = n; k0 += W0; k1 += W1;
    }
    return c;
}

// four N(0,1) draws of counter (row, quad, step): u=((x>>9)+0.5)*2^-23, Box-Muller (cos,sin)
__device__ __forceinline__ void philox_normal4(uint64_t seed, uint64_t row, uint32_t quad, uint32_t step, float n[4]) {
    u32x4 c{(uint32_t)row, (uint32_t)(row >> 32), quad, step};
    u32x4 r = philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    const float s = 1.1920928955078125e-07f;      // 2^-23
    float u0 = ((float)(r.x >> 9) + 0.5f) * s, u1 = ((float)(r.y >> 9) + 0.5f) * s;
    float u2 = ((float)(r.z >> 9) + 0.5f) * s, u3 = ((float)(r.w >> 9) + 0.5f) * s;
    const float twopi = 6.2831854820251465f;       // fp32(2*pi)
    float r0 = sqrtf(-2.f * logf(u0)), r1 = sqrtf(-2.f * logf(u2));
    float s0, c0, s1, c1;
    sincosf(twopi * u1, &s0, &c0);
    sincosf(twopi * u3, &s1, &c1);
    n[0] = r0 * c0; n[1] = r0 * s0; n[2] = r1 * c1; n[3] = r1 * s1;
}

}  // namespace oevae
