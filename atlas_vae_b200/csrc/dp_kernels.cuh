// Data-parallel update over NVLink peer memory: ONE kernel replaces "sum the partial slices" +
// ncclAllReduce(SUM) + "clip_values + Adam" of the NCCL path.
//   phase 1  block b sums this rank's partial-gradient slices for its 256 parameters (fixed order, adam_kernels.cuh) and stores the result
//            in the rank's exchange buffer, then pushes sequence number `seq` into flag [rank][b] of EVERY peer (P2P store);
//   phase 2  it waits until flags [q][b] of all peers q have reached `seq` (they are written into LOCAL memory by the
//            peers), reads the same 256 values from every peer's exchange buffer (P2P loads), adds the ranks in rank
//            order -- every rank computes bit-identical sums, so the replicas stay bit-identical -- and applies
//            clip_values (models.py:185) + Keras Adam (models.py:138,186) + the bf16 weight-image refresh.
// Nothing is grid-wide: a block depends only on block b of each peer.  All blocks of the launch must be co-resident
// (a waiting block must never keep a producing block of this rank from being scheduled); the host checks that.
// The exchange buffer is double-buffered by step parity: a peer overwrites buffer p of step k+2 only after its own
// step k+1 update, which waited for this rank's step k+1 flags, which this rank raised after its step k kernel
// (and with it every read of buffer p) had completed.
#pragma once
#include "common.cuh"
#include "umma.cuh"
#include "adam_kernels.cuh"

namespace oevae {

constexpr int kDpMaxWorld = 8;               // blocks: ceil(P / kAdamCols) of kAdamThreads threads -- 287 at cfg-2, one co-resident wave

struct DpArgs {
    const float* parts; int n_parts; long part_stride;
    float* my_gred;                          // this rank's exchange buffer of the step (local)
    const float* peer_gred[kDpMaxWorld];     // the same buffer of every rank (own entry = my_gred)
    unsigned int* my_flags;                  // [world][gridDim.x], written by the peers
    unsigned int* peer_flags[kDpMaxWorld];   // every rank's flag array (own entry unused)
    int rank, world; unsigned int seq;
    float* theta; float* m; float* v; long n;
    float alpha, b1, b2, eps;
    unsigned char* wimg; const int* pos;
    float* gsum_out;                         // nullable [n]: SUM over ranks before clip_values (parity checks)
    long long* stamps;                       // nullable [gridDim.x][6] globaltimer ns (OEVAE_DP_STAMPS=1): entry, dependency met, local sum published,
                                             // all peer flags seen, peer loads done, exit
};

__global__ void __launch_bounds__(kAdamThreads, 2) dp_adam_kernel(DpArgs a) {
    __shared__ __align__(16) float red[kAdamSlices][kAdamCols];
    // PDL: the next step's training kernel may begin its set-up (barriers, TMEM, zeroed tiles, X prefetch) on the SMs this kernel's
    // blocks leave; it orders itself behind this grid with griddepcontrol.wait.  Dependents are released only once EVERY block of
    // this grid has started, so a waiting block can never be kept off the chip by them.
    auto stamp = [&](int k) { if (a.stamps && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); a.stamps[6 * blockIdx.x + k] = (long long)t; } };
    stamp(0);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");      // the training kernel's partial slices are complete
    stamp(1);
    const long i = (long)blockIdx.x * kAdamCols + threadIdx.x;
    const bool in = threadIdx.x < kAdamCols && i < a.n;
    // the update's own operands do not depend on the exchange: their loads fly during phase 1
    float th0 = 0.f, m0 = 0.f, v0 = 0.f;
    int p0 = -1, p1 = -1, p2 = -1, p3 = -1;
    if (in) {
        th0 = a.theta[i]; m0 = a.m[i]; v0 = a.v[i];
        if (a.wimg) { p0 = a.pos[i]; p1 = a.pos[a.n + i]; p2 = a.pos[2 * a.n + i]; p3 = a.pos[3 * a.n + i]; }
    }
    const float g = reduce_parts(a.parts, a.n_parts, a.part_stride, red);
    if (in) __stcg(a.my_gred + i, g);
    __syncthreads();
    // ---- publish: this block's sums are visible system-wide before any peer sees the flag
    // (st.release.sys is cumulative over the block's stores ordered before it by the barrier above)
    if (threadIdx.x < a.world && (int)threadIdx.x != a.rank) {
        unsigned int* f = a.peer_flags[threadIdx.x] + (long)a.rank * gridDim.x + blockIdx.x;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(a.seq) : "memory");
    }
    stamp(2);
    // ---- wait for block blockIdx.x of every peer
    if (threadIdx.x < a.world && (int)threadIdx.x != a.rank) {
        const unsigned int* f = a.my_flags + (long)threadIdx.x * gridDim.x + blockIdx.x;
        unsigned int seen;
        const long long t0 = clock64();
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
            if ((int)(seen - a.seq) < 0 && clock64() - t0 > 20000000000ll) __trap();   // ~10 s: a peer died; fail loudly instead of hanging
        } while ((int)(seen - a.seq) < 0);
    }
    __syncthreads();
    stamp(3);
    if (!in) { return; }
    float tq[kDpMaxWorld];
#pragma unroll
    for (int q = 0; q < kDpMaxWorld; ++q) {                  // all peers' loads in flight together
        tq[q] = 0.f;
        if (q < a.world && q != a.rank) asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(tq[q]) : "l"(a.peer_gred[q] + i) : "memory");
    }
    float gs = 0.f;
#pragma unroll
    for (int q = 0; q < kDpMaxWorld; ++q) if (q < a.world) gs += (q == a.rank) ? g : tq[q];
    stamp(4);
    if (a.gsum_out) a.gsum_out[i] = gs;
    gs = clipv(gs);
    float m = m0, v = v0;
    m += (gs - m) * (1.f - a.b1);
    v += (gs * gs - v) * (1.f - a.b2);
    a.m[i] = m; a.v[i] = v;
    const float th = th0 - (m * a.alpha) / (sqrtf(v) + a.eps);
    a.theta[i] = th;
    if (a.wimg) {
        __nv_bfloat16 hi, lo;
        umma::split_bf16(th, hi, lo);
        if (p0 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p0) = hi;
        if (p1 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p1) = lo;
        if (p2 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p2) = hi;
        if (p3 >= 0) *reinterpret_cast<__nv_bfloat16*>(a.wimg + p3) = lo;
    }
    stamp(5);
}

}  // namespace oevae
