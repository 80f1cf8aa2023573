// Data-parallel update over NVLink peer memory: ONE kernel replaces "sum the partial slices" +
// ncclAllReduce(SUM) + "clip_values + Adam" of the NCCL path.
//   phase 1  block b sums this rank's partial-gradient slices for its 256 parameters (fixed order, adam_kernels.cuh) and PUSHES
//            the sums into its 86 slots [rank][86 b + j] of every peer's receive buffer, three sums and the step's tag per slot, ONE
//            16-byte store {v0, v1, v2, seq} each (P2P store);
//   phase 2  every thread polls the slot of its parameter in its OWN receive buffer (local memory), once per peer, until the tag
//            is this step's `seq`, adds the ranks in rank order -- every rank computes bit-identical sums, so the replicas stay
//            bit-identical -- and applies clip_values (models.py:185) + Keras Adam (models.py:138,186) + the weight-image refresh.
// An aligned 16-byte vector store is one transaction, so values and tag arrive together: no flag, no fence, no barrier, one
// NVLink traversal.  At 8 GPUs the exchange is bound by the bytes a GPU pushes (7 peers x slots): 8-byte {value, tag} slots were
// 4.1 MB per GPU and step (5.6-9 us wait, profiles/r02_dp_stamps_n8_ll8.txt), 16-byte slots are 2.8 MB.  (Round 1 published a per-block flag with st.release.sys after storing the sums locally and the peers pulled them:
// the 287 system-scope releases alone cost a median 4.4 us per block, 8 us for the slowest, then flag wait + a load round trip --
// profiles/r02_dp_stamps_n2_pull.txt.)
// Nothing is grid-wide: a thread depends only on the same parameter of each peer.  All blocks of the launch must be
// co-resident (a waiting block must never keep a producing block of this rank from being scheduled); the host checks that.
// The receive buffer is double-buffered by step parity: a peer overwrites parity p again at step k+2, after its own step
// k+1 update, which needed this rank's step k+1 sums, which this rank pushed after its step k kernel (and with it every
// read of parity p) had completed.
#pragma once
#include "common.cuh"
#include "umma.cuh"
#include "adam_kernels.cuh"

namespace oevae {

constexpr int kDpStamps = 8;                 // globaltimer stamps per block (OEVAE_DP_STAMPS=1)
constexpr int kDpSlots = (kAdamCols + 2) / 3;  // 16-byte slots per block: three sums + the tag
constexpr int kDpMaxWorld = 8;               // blocks: ceil(P / kAdamCols) of kAdamThreads threads -- 287 at cfg-2, one co-resident wave

struct DpArgs {
    EarlyGroups early;                       // adam_kernels.cuh: blocks of early parameter groups start before the training grid ends
    SumsTrace trace;                         // adam_kernels.cuh: the step's running loss sums -> mapped host memory
    const float* parts; int n_parts; long part_stride;
    const uint4* my_recv;                    // this rank's receive buffer of the step's parity: [world][stride] {v0, v1, v2, seq}, written by the peers
    uint4* peer_recv[kDpMaxWorld];           // the same buffer of every rank (own entry unused)
    long stride;                             // slots per rank in a receive buffer (>= gridDim.x * kDpSlots)
    int rank, world; unsigned int seq;       // seq: this step's tag, never 0 (the buffers start zeroed)
    float* theta; float* m; float* v; long n;
    float alpha, b1, b2, eps;
    unsigned char* wimg; const int* pos;
    float* gsum_out;                         // nullable [n]: SUM over ranks before clip_values (parity checks)
    long long* stamps;                       // nullable [gridDim.x][kDpStamps] globaltimer ns (OEVAE_DP_STAMPS=1): entry, dependency met, local sum ready,
                                             // sums pushed, all peer sums seen, exit
};

__global__ void __launch_bounds__(kAdamThreads, 2) dp_adam_kernel(DpArgs a) {
    __shared__ __align__(16) float red[kAdamSlices][kAdamCols];
    // PDL: the next step's training kernel may begin its set-up (barriers, TMEM, zeroed tiles, X prefetch) on the SMs this kernel's
    // blocks leave; it orders itself behind this grid with griddepcontrol.wait.  Dependents are released only once EVERY block of
    // this grid has started, so a waiting block can never be kept off the chip by them.
    auto stamp = [&](int k) { if (a.stamps && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); a.stamps[kDpStamps * blockIdx.x + k] = (long long)t; } };
    stamp(0);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int chunk = (int)gridDim.x - 1 - (int)blockIdx.x;  // reversed: the first-resident blocks own the output layer (EarlyGroups)
    const long c0 = (long)chunk * kAdamCols;
    // the training kernel's partial slices are complete -- for the whole grid, or for this block's early parameter group
    if (!early_wait(a.early, c0, min(c0 + kAdamCols, a.n))) asm volatile("griddepcontrol.wait;" ::: "memory");
    stamp(1);
    const long i = c0 + threadIdx.x;
    const bool in = threadIdx.x < kAdamCols && i < a.n;
    // the update's own operands do not depend on the exchange: their loads fly during phase 1 (issued only now: an early block may
    // be resident while the PREVIOUS update kernel is still writing them)
    float th0 = 0.f, m0 = 0.f, v0 = 0.f;
    int p0 = -1, p1 = -1, p2 = -1, p3 = -1;
    if (in) {
        th0 = a.theta[i]; m0 = a.m[i]; v0 = a.v[i];
        if (a.wimg) { p0 = a.pos[i]; p1 = a.pos[a.n + i]; p2 = a.pos[2 * a.n + i]; p3 = a.pos[3 * a.n + i]; }
    }
    const float g = reduce_parts(a.parts, a.n_parts, a.part_stride, red, chunk);
    stamp(2);
    write_trace(a.trace, chunk);
    // ---- push {three sums, seq} to every peer: thread j < 86 sends slot j of the block (a warp writes 512 contiguous bytes per peer)
    __syncthreads();                                         // red[][] has been read by every thread
    if (threadIdx.x < kAdamCols) red[0][threadIdx.x] = g;
    __syncthreads();
    if (threadIdx.x < kDpSlots) {
        const int j = 3 * threadIdx.x;
        const uint32_t v0 = __float_as_uint(red[0][j]), v1 = j + 1 < kAdamCols ? __float_as_uint(red[0][j + 1]) : 0u, v2 = j + 2 < kAdamCols ? __float_as_uint(red[0][j + 2]) : 0u;
        const long slot = (long)a.rank * a.stride + (long)chunk * kDpSlots + threadIdx.x;
#pragma unroll
        for (int q = 0; q < kDpMaxWorld; ++q)
            if (q < a.world && q != a.rank)
                asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(a.peer_recv[q] + slot), "r"(v0), "r"(v1), "r"(v2), "r"(a.seq) : "memory");
    }
    stamp(3);
    if (!in) { return; }
    // ---- collect the peers' sums of this parameter from the local receive buffer
    float tq[kDpMaxWorld];
    unsigned int pending = 0;
#pragma unroll
    for (int q = 0; q < kDpMaxWorld; ++q) { tq[q] = 0.f; if (q < a.world && q != a.rank) pending |= 1u << q; }
    const long my_slot = (long)chunk * kDpSlots + threadIdx.x / 3;
    const int comp = threadIdx.x % 3;
    const long long t0 = clock64();
    while (pending) {
#pragma unroll
        for (int q = 0; q < kDpMaxWorld; ++q) {
            if (pending & (1u << q)) {
                uint32_t r0, r1, r2, tag;
                asm volatile("ld.relaxed.sys.global.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(tag) : "l"(a.my_recv + (long)q * a.stride + my_slot) : "memory");
                if (tag == a.seq) { tq[q] = __uint_as_float(comp == 0 ? r0 : (comp == 1 ? r1 : r2)); pending &= ~(1u << q); }
            }
        }
        if (pending && clock64() - t0 > 20000000000ll) __trap();   // ~10 s: a peer died; fail loudly instead of hanging
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
