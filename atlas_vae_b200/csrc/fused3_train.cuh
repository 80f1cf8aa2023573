// Warp-specialised version of the feature-major fused OE-VAE step (same plan, shared-memory layout, weight image and
// arithmetic as fused2_train.cuh; see that file and fused2_plan.h for the layouts).
//   forward + losses  models.py:30-36,49-52,62-69,72-85,123-132      gradient of the batch SUM  models.py:183
//   scoring (SCORE)   vae.py:163-167, utils.py:346-386
// Roles (704 threads):
//   warps 0..15  workers: stage X chunks (fp32 -> bf16 hi/lo), run every epilogue (bias/ReLU/split, Sampling + KLD, loss and
//                its gradient, ReLU masks)
//   warp 16      MMA issuer (one lane): waits for operands on mbarriers, issues tcgen05.mma, signals with tcgen05.commit
//   warp 17      weight producer (one lane): cp.async.bulk of the pre-swizzled weight ops through a 4-slot ring
//   warps 18..21 drain warps (one per TMEM lane quarter): dW^T accumulators + bias-gradient parts -> partial-gradient slice.
//                The slice is 293 KB per tile at cfg-2 and one SM stores ~17 B/clk, so the drains run beside the chain
// SCORE instantiations (forward only, 640 threads): warps 18..19 are X stagers instead (X chunks of the NEXT tile -> X ring), the
//   weight ops follow the SCORE schedule (layer 0 of tile t+1 interleaved with the chain of tile t), tiles come from an atomic
//   counter through a shared-memory queue, decoder tiles alias encoder tiles, layer 0 accumulates in TMEM columns [320, 448).
// No __syncthreads inside a tile: every hand-off is an mbarrier, so the issuer runs ahead of the epilogues, X staging
// overlaps the layer-0 MMAs, and the dW^T drains overlap the next stage's MMAs (two dW^T accumulators).
//
// TMEM columns: [0,64) mu, [64,128) log_var, [128,192) chain accumulator, [192,256) / [256,320) output-layer chunk
// accumulators (slot 0 doubles as the second layer-0 accumulator), [320,512) dW^T accumulators (one or two).
#pragma once
#include "fused2_train.cuh"

namespace oevae {

constexpr int kF3Workers = 16;                               // worker warps
constexpr int kF3Drains = 4;                                 // drain warps, one per TMEM lane quarter
constexpr int kF3Threads = (kF3Workers + 2 + kF3Drains) * 32;
constexpr int kF3XSlotsMax = 4;                              // SCORE: X ring slots (2 in the plan's X region + 2 over the aliased decoder tiles)
constexpr int kF3Stagers = 2;                                // SCORE: X stager warps (run one tile ahead of the workers)
constexpr int kF3ScoreThreads = (kF3Workers + 2 + kF3Stagers) * 32;
constexpr int kF3MaxChunks = 24;                             // 128-feature chunks of the output layer (plan: 2 nX + 4 n + 8 <= 96 ops)
constexpr uint32_t kF3_TDW = 320;

namespace f3 {
using namespace umma;
// mma_e / commit_e / arrive_e (umma.cuh): single-thread tcgen05 instructions issued from warp-uniform code
// D[128 x N] (+)= A * B over nk k-steps, bf16x3 (Ahi*Bhi + Ahi*Blo + Alo*Bhi); see f2::issue_core
__device__ __forceinline__ void issue_e(uint32_t d, const f2::Opnd& A, const f2::Opnd& B, int nk, int N, bool acc, int M = 128) {
    uint64_t ah = f2::desc(A, 0, false), al = f2::desc(A, 0, true), bh = f2::desc(B, 0, false), bl = f2::desc(B, 0, true);
    const uint32_t a_step = A.kstep >> 4, b_step = B.kstep >> 4, idesc = idesc_bf16_major(M, N, A.mn, B.mn);
    uint32_t ac = acc ? 1u : 0u;
#pragma unroll 1
    for (int ks = 0; ks < nk; ++ks) {
        mma_e(d, ah, bh, idesc, ac);
        mma_e(d, ah, bl, idesc, 1u);
        mma_e(d, al, bh, idesc, 1u);
        ac = 1u;
        ah += a_step; al += a_step; bh += b_step; bl += b_step;
    }
}
// same, k-steps alternating between two accumulators (d0 even, d1 odd); the epilogue adds the two
__device__ __forceinline__ void issue2_e(uint32_t d0, uint32_t d1, const f2::Opnd& A, const f2::Opnd& B, int nk, int N, bool acc0, bool acc1) {
    uint64_t ah = f2::desc(A, 0, false), al = f2::desc(A, 0, true), bh = f2::desc(B, 0, false), bl = f2::desc(B, 0, true);
    const uint32_t a_step = A.kstep >> 4, b_step = B.kstep >> 4, idesc = idesc_bf16_major(128, N, A.mn, B.mn);
    uint32_t ac0 = acc0 ? 1u : 0u, ac1 = acc1 ? 1u : 0u;
#pragma unroll 1
    for (int ks = 0; ks < nk; ++ks) {
        const uint32_t d = (ks & 1) ? d1 : d0, ac = (ks & 1) ? ac1 : ac0;
        mma_e(d, ah, bh, idesc, ac);
        mma_e(d, ah, bl, idesc, 1u);
        mma_e(d, al, bh, idesc, 1u);
        if (ks & 1) ac1 = 1u; else ac0 = 1u;
        ah += a_step; al += a_step; bh += b_step; bl += b_step;
    }
}
// One element of utils.py:346-375's KLD and JSD sums for tame operands (see the caller): kl += nan_to_num(p log2(p/q)),
// js += nan_to_num(p log2(p/m)) + nan_to_num(q log2(q/m)) with m = (p + q) / 2 -- the caller halves the JSD sum once at the end
// (exact in binary).  With a = p/m, b = q/m (one SFU reciprocal, a + b = 2): log2(p/q) = log2 a - log2 b.  A term whose ratio is
// not positive is NaN (sign differs) or 0 * -inf (p = 0) in NumPy, which nan_to_num turns into 0: those terms are skipped.
__device__ __forceinline__ void score_kj(float p, float q, float& kl, float& js) {
    const float m = (p + q) * 0.5f;
    // one block, so that ptxas keeps the three accumulations as predicated FFMAs (as C++ it turned each element into a branch)
    asm("{\n\t.reg .pred pa, pb, pc;\n\t.reg .f32 rm, av, bv, la, lb, dl;\n\t"
        "rcp.approx.ftz.f32 rm, %4;\n\t"
        "mul.f32 av, %2, rm;\n\t"
        "mul.f32 bv, %3, rm;\n\t"
        "lg2.approx.ftz.f32 la, av;\n\t"
        "lg2.approx.ftz.f32 lb, bv;\n\t"
        "setp.gt.f32 pa, av, 0f00000000;\n\t"
        "setp.gt.f32 pb, bv, 0f00000000;\n\t"
        "and.pred pc, pa, pb;\n\t"
        "sub.f32 dl, la, lb;\n\t"
        "@pa fma.rn.f32 %1, %2, la, %1;\n\t"
        "@pb fma.rn.f32 %1, %3, lb, %1;\n\t"
        "@pc fma.rn.f32 %0, %2, dl, %0;\n\t}"
        : "+f"(kl), "+f"(js) : "f"(p), "f"(q), "f"(m));
}
__device__ __forceinline__ void score_kj_slow(float p, float q, float& kl, float& js) {
    const float m = (p + q) * 0.5f;
    kl += f2::kld_term_fast(p, q);
    js += f2::kld_term_fast(p, m) + f2::kld_term_fast(q, m);              // halved by the caller
}
// TMEM -> registers: 32 lanes x 4 consecutive columns
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void worker_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kF3Workers * 32) : "memory"); }
// every lane has fenced its own writes / TMEM reads; one arrival per warp
__device__ __forceinline__ void warp_arrive(uint64_t* bar, int lane) {
    __syncwarp();
    if (lane == 0) mbar_arrive(bar);
}
}  // namespace f3

// MODE 0: one input (training step, get_losses, scoring).  MODE 1: paired tiles (OE_type 'KLD').  MODE 2: two inputs [bkg ; OoD] with
// per-jet loss-term coefficients (raw pre-pass and training pass of the reconstruction-based OE types).
template <bool SCORE, int MODE = 0>
__global__ void __launch_bounds__(SCORE ? kF3ScoreThreads : kF3Threads, 1) fused3_kernel(const __grid_constant__ F2Plan P, F2Args a) {
    using namespace umma;
    using namespace f2;
    using namespace f3;
    constexpr bool PAIRED = MODE == 1, TWOIN = MODE == 2;
    constexpr bool NOMSE = SCORE && MODE == 3;               // scoring pass without the MSE vector: 16 fewer live accumulators
    // Second layer-0 accumulator.  SCORE: the X stager warps and the issuer run layer 0 of tile t+1 while the workers are still in the
    // output epilogues of tile t, so it must not be a y-chunk accumulator; the dW^T columns are free in a forward pass.
    constexpr uint32_t T_L0A = SCORE ? kF3_TDW : f2::T_CH, T_L0B = SCORE ? kF3_TDW + 64 : f2::T_Y0;
    const bool is_train = !SCORE && a.train;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t wfull[kF2WSlots], wempty[kF2WSlots], xfull[kF3XSlotsMax], xempty[kF3XSlotsMax];
    __shared__ uint64_t yfull[2], yempty[2], gyfull[kF3MaxChunks], gyempty, dwfull[2], dwempty[2];
    __shared__ uint64_t acc[kF2MaxStages], tready[kF2MaxStages], accB[kF2MaxStages], gready[kF2MaxStages];
    __shared__ uint32_t tmem_base_s;
    __shared__ int tile_q[16], tile_seq[16];                 // SCORE: dynamic tile queue (iteration i -> slot i & 15), see tile_of
    unsigned char* sm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler: role branches do not diverge
    const int q = warp & 3, h = (warp >> 2) & 3;
    const int frow = 32 * q + lane;                          // TMEM lane = feature row owned by this worker thread
    const int jc0 = 16 * h;                                  // first of the 16 jet columns owned by this worker thread
    const int n = P.n, D = P.D, L = P.L, so = 2 * n + 1;
    const int nops = is_train ? P.n_op_train : P.n_op_eval;
    const int n_fwd_ops = P.op_first[so] + P.nX;
    const int ndw = P.n_dwbuf, dwc = P.dw_cols;             // dW^T accumulators and their column pitch

    float* s_w = reinterpret_cast<float*>(sm + P.scratch_off);        // [64] jet weights (0 for rows past the batch)
    float* s_rec = s_w + 64;                                           // [4][64] per lane-quarter reconstruction sums
    float* s_kld = s_rec + 256;                                        // [64]
    float* s_db = s_kld + 64;                                          // [6 buffers][4 column groups][128] bias-gradient parts:
                                                                       // chain stage k -> buffer k % 3, output chunk c -> 3 + c % 3 (a buffer
                                                                       // is rewritten two worker_sync()s after its last reader)
    float* s_eps = s_db + 3072;                                        // [L][64]
    float* s_wk = s_eps + 64 * L;                                      // [64] per-jet coefficient of the KLD term (beta * w, or c_kld)
    // Tiles of an outlier-exposure batch (a.x2 != null): the bkg rows [0, B1) fill tiles 0 .. ceil(B1 / 64) - 1 and the OoD rows
    // [B1, B) start at the NEXT tile, so a tile never mixes the two inputs.  tile_src: first row of the tile in its input, the
    // input pointer at that row, and how many of the 64 jets exist.
    // Paired mode (PAIRED, OE_type 'KLD', models.py:88-96): tile t holds pairs [32 t, 32 t + 32) -- jet columns 0..31 are the bkg
    // rows, columns 32..63 their OoD partners -- so relu(KLD(bkg) - KLD(OoD) + margin) and its derivative are formed inside the
    // tile right after the heads and the whole step is one launch.  tile_src then returns the bkg pointer of the tile.
    auto tile_src = [&](int tile, int& local0, int& nvalid) -> const float* {          // row counts fit 31 bits (oevae.h)
        if (PAIRED) {
            local0 = tile * 32; nvalid = max(0, min(32, (int)a.B - local0));
            return a.x + (long)local0 * P.D;
        }
        if (!TWOIN) {
            local0 = tile * kF2Jets; nvalid = max(0, min(kF2Jets, (int)a.B - local0));
            return a.x + (long)local0 * P.D;
        }
        const int t1 = (int)((a.B1 + kF2Jets - 1) / kF2Jets);                             // first tile of the second input
        const bool second = tile >= t1;
        local0 = (second ? tile - t1 : tile) * kF2Jets;
        nvalid = max(0, min(kF2Jets, (int)(second ? a.B - a.B1 : a.B1) - local0));
        return (second ? a.x2 : a.x) + (long)local0 * P.D;
    };
    // SCORE (forward only): decoder tile k reuses the bytes of encoder tile 2n - k (same rows in the mirrored decoder, and H_{2n-k}
    // is dead once stage 2n-k+1 has run), and the bytes of tiles n+1..2n hold two more X ring slots when they fit, so the stagers
    // run up to four chunks ahead of the layer-0 MMAs.
    const bool alias_ok = SCORE && [&] { for (int k = n + 1; k <= 2 * n; ++k) if (P.tile_rows[k] != P.tile_rows[2 * n - k]) return false; return n >= 1; }();
    const int xs_log = (alias_ok && P.wring_off - P.tile_off[n + 1] >= 32768) ? 2 : 1;       // log2(X ring slots)
    const uint32_t xs_mask = (1u << xs_log) - 1u;
    auto toff = [&](int k) { return P.tile_off[(alias_ok && k > n) ? 2 * n - k : k]; };
    auto tile_hi = [&](int k) { return sm + toff(k); };
    auto tile_lo = [&](int k) { return sm + toff(k) + P.tile_rows[k] * 128; };
    auto xslot_off = [&](int s) { return s < 2 ? P.s_off + s * 16384 : P.tile_off[n + 1] + (s - 2) * 16384; };
    auto xslot = [&](int s) { return sm + xslot_off(s); };             // hi 8 KB, lo 8 KB
    unsigned char* gy_hi = sm + P.s_off; unsigned char* gy_lo = gy_hi + 16384;
    auto db_at = [&](int buf, int hh, int row) -> float& { return s_db[(buf * 4 + hh) * 128 + row]; };

    unsigned long long dbg_t0 = 0;
    if (a.dbg && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
    long long dbg_c[6] = {0, 0, 0, 0, 0, 0};
    dbg_c[0] = clock64();
    // the first tile's rows of X start moving towards L2 while the CTA sets itself up
    if (warp < kF3Workers && (int)blockIdx.x < a.n_tiles) {
        int l0, nv;
        const char* px = reinterpret_cast<const char*>(tile_src((int)blockIdx.x, l0, nv));
        const long nbytes = (long)nv * P.D * 4;
        for (long o = (long)tid * 128; o < nbytes; o += kF3Workers * 32 * 128) prefetch_l2(px + o);
    }
    dbg_c[1] = clock64();
    // ---- one-time setup
    // barrier init spread over the threads (one thread doing all ~110 of them costs ~1.3 k cycles)
    if (tid < kF2MaxStages) { mbar_init(&acc[tid], 1); mbar_init(&tready[tid], kF3Workers); mbar_init(&accB[tid], 1); mbar_init(&gready[tid], kF3Workers); }
    else if (tid >= 32 && tid < 32 + kF3MaxChunks) mbar_init(&gyfull[tid - 32], kF3Workers);
    else if (tid >= 64 && tid < 64 + kF2WSlots) { mbar_init(&wfull[tid - 64], 1); mbar_init(&wempty[tid - 64], 1); }
    else if (tid >= 96 && tid < 96 + kF3XSlotsMax) { mbar_init(&xfull[tid - 96], SCORE ? kF3Stagers : kF3Workers); mbar_init(&xempty[tid - 96], 1); }
    else if (tid >= 128 && tid < 130) { mbar_init(&yfull[tid - 128], 1); mbar_init(&yempty[tid - 128], kF3Workers); mbar_init(&dwfull[tid - 128], 1); mbar_init(&dwempty[tid - 128], kF3Drains); }
    else if (tid == 160) mbar_init(&gyempty, 1);
    if (tid >= 192 && tid < 208) tile_seq[tid - 192] = -1;
    mbar_fence_init();
    dbg_c[2] = clock64();
    for (int i = tid * 16; i < P.zero_bytes; i += (int)blockDim.x * 16) *reinterpret_cast<uint4*>(sm + i) = make_uint4(0, 0, 0, 0);
    dbg_c[3] = clock64();
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    dbg_c[4] = clock64();
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    dbg_c[5] = clock64();
    if (a.dbg && blockIdx.x == 0 && tid == 0) for (int i = 0; i < 6; ++i) a.dbg[900 + i] = dbg_c[i];
    // Programmatic dependent launch: everything above (X prefetch, barriers, TMEM, zeroed tiles) overlaps the tail of the
    // previous kernel in the stream (the Adam / weight-image kernel); its results (weight image, biases) and the partial-
    // gradient slices it reads are only touched below.  Both instructions are no-ops for an ordinary launch.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // Training / get_losses: the worker warps first stage the first two X chunks of their tile (inputs only: nothing the previous
    // kernel writes) and execute their griddepcontrol.wait after that, so ~3 k cycles of layer 0 also overlap the update kernel.
    // (not MODE 2: its per-jet coefficients c_rec / c_kld are written by the combine kernel right before this launch)
    const bool late_wait = !SCORE && MODE != 2 && warp < kF3Workers;
    if (!late_wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t tm = tmem_base_s;
    const uint32_t t_lane = tm + ((uint32_t)(32 * q) << 16);
    const int my_tiles = ((int)blockIdx.x < a.n_tiles) ? (a.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    int n_dbg = 0;
    bool dbg_on = true;                                      // the workers switch the stamps on from their tile a.dbg[1000] on
    const int dbg_skip = a.dbg ? (int)a.dbg[1000] : 0;
    auto stamp = [&]() { if (a.dbg && dbg_on && blockIdx.x == 0 && tid == 0 && n_dbg < 255) a.dbg[1 + n_dbg++] = clock64(); };
    stamp();
    int n_dbg2 = 0;                                          // MMA issuer's own stamps: a.dbg[256] = count, [257..511]
    auto stamp2 = [&]() { if (a.dbg && blockIdx.x == 0 && n_dbg2 < 255) a.dbg[257 + n_dbg2++] = clock64(); };

    // SCORE tile schedule.  SMs do not run at the same speed (measured: the slowest CTA of a 100-tile pass needs 8 % longer than the
    // median, single TPCs up to 30 %), so tiles beyond the first are handed out by an atomic counter instead of round-robin.  Lane 0 of
    // the first stager warp (the role that runs furthest ahead) publishes the tile of iteration i + 2 when it starts iteration i; every
    // role reads the tile of its iteration from the queue (-1 ends the pass).  No role can be more than 5 iterations behind the stagers
    // (the X ring holds at most 4 tiles -- one-chunk inputs -- and the acc[0] phase bounds the issuer), so 16 slots need no hand-back.
    auto tile_pub = [&](int it) {
        int id = it == 0 ? (int)blockIdx.x : (int)gridDim.x + atomicAdd(a.tile_ctr, 1);
        if (id >= a.n_tiles) id = -1;
        tile_q[it & 15] = id;
        __threadfence_block();
        *reinterpret_cast<volatile int*>(&tile_seq[it & 15]) = it;
        return id;
    };
    auto tile_of = [&](int it) -> int {
        if (!SCORE || !a.tile_ctr) { const long t_ = (long)blockIdx.x + (long)it * gridDim.x; return t_ < a.n_tiles ? (int)t_ : -1; }
        else {
            while (*reinterpret_cast<volatile int*>(&tile_seq[it & 15]) != it) __nanosleep(32);
            __threadfence_block();
            return *reinterpret_cast<volatile int*>(&tile_q[it & 15]);
        }
    };

    // SCORE weight-op schedule.  The ring is consumed in order, so the order of the ops IS the issuer's schedule: layer 0 of tile
    // t+1 is interleaved with the chain of tile t (the issuer is idle while the workers run a chain epilogue, the stagers have
    // the X chunks ready, and the layer-0 accumulators [320, 448) are nobody else's in a forward pass).  Producer and issuer walk
    // the same sequence:
    //   prologue     L0 chunks 0..nX-1 of the CTA's first tile, dummies up to a multiple of the ring
    //   per tile     for s = 1..2n: chain op s, then sc_cnt(s) L0' chunks (nX spread over the stages, the remainder from the heads
    //                on: their epilogue is the longest); a dummy if the position is odd (an output chunk reads two adjacent
    //                slots); the nX output ops; dummies up to a multiple of the ring
    // L0' = chunk of the NEXT tile of this CTA (a 16-byte dummy after the last tile).
    const int sc_n2 = 2 * n;
    auto sc_cnt = [&](int st_) { return P.nX / sc_n2 + (((st_ - n + sc_n2) % sc_n2) < P.nX % sc_n2 ? 1 : 0); };

    if (SCORE && warp == kF3Workers + 1) {
        // ======================================================================================= weight producer (SCORE schedule)
        if (lane == 0) {
            long g = 0;
            auto put = [&](int k) {                                          // plan op k, or -1 = 16-byte dummy
                const int slot = (int)(g & (kF2WSlots - 1));
                if (g >= kF2WSlots) mbar_wait(&wempty[slot], (uint32_t)(((g >> 2) - 1) & 1));
                const uint32_t bytes = k >= 0 ? P.op_bytes[k] : 16u;
                mbar_expect_tx(&wfull[slot], bytes);
                bulk_g2s(sm + P.wring_off + slot * P.wslot_bytes, a.wimg + (k >= 0 ? P.op_off[k] : 0u), bytes, &wfull[slot]);
                ++g;
            };
            for (int c = 0; c < P.nX; ++c) put(P.op_first[0] + c);
            while (g & (kF2WSlots - 1)) put(-1);
            for (int it = 0; tile_of(it) >= 0; ++it) {
                const bool last = tile_of(it + 1) < 0;
                int c = 0;
                for (int st_ = 1; st_ <= sc_n2; ++st_) {
                    put(P.op_first[st_]);
                    for (int i = sc_cnt(st_); i > 0; --i, ++c) put(last ? -1 : P.op_first[0] + c);
                }
                if (g & 1) put(-1);
                for (int c2 = 0; c2 < P.nX; ++c2) put(P.op_first[so] + c2);
                while (g & (kF2WSlots - 1)) put(-1);
            }
        }
        __syncwarp();
    } else if (SCORE && warp == kF3Workers) {
        // ======================================================================================= MMA issuer (SCORE schedule)
        // every lane runs this code with identical values (see f3::mma_e)
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base_s, 0);
        const uint32_t smu = __shfl_sync(0xffffffffu, smem_u32(sm), 0);
        auto tile_a = [&](int k) { return smu + (uint32_t)toff(k); };
        auto xslot_a = [&](uint32_t b) { return smu + (uint32_t)xslot_off((int)b); };
        long wseen = 0;                                                      // ring positions [0, wseen) have landed
        long wp = 0;                                                         // next ring position to consume
        auto slot_a = [&](long p_) { return smu + (uint32_t)P.wring_off + (uint32_t)(p_ & (kF2WSlots - 1)) * (uint32_t)P.wslot_bytes; };
        auto wgetp = [&](long p_) -> uint32_t { for (; wseen <= p_; ++wseen) mbar_wait(&wfull[wseen & (kF2WSlots - 1)], (uint32_t)((wseen >> 2) & 1)); return slot_a(p_); };
        auto wrelp = [&](long p_) { commit_e(&wempty[p_ & (kF2WSlots - 1)]); };
        auto wskipp = [&](long p_) { wgetp(p_); arrive_e(&wempty[p_ & (kF2WSlots - 1)]); };
        uint32_t xk = 0, yk = 0;
        auto issue_l0 = [&](long p_, int c) {                                // chunk c of a layer 0, weights at ring position p_
            const F2Stage& st = P.st[0];
            const uint32_t wb = wgetp(p_);
            const uint32_t xb_ = xk & xs_mask;
            mbar_wait(&xfull[xb_], (xk >> xs_log) & 1);
            tc_fence_after();
            const int nk = min(4, (D - 64 * c + 15) >> 4);
            issue2_e(tm + T_L0A, tm + T_L0B, opnd_k(wb, st.rows * 128), opnd_k(xslot_a(xb_), 8192), nk, kF2Jets, c > 0, c > 0);
            commit_e(&xempty[xb_]);
            wrelp(p_);
            ++xk;
            if (c == P.nX - 1) commit_e(&acc[0]);
        };
        {
            for (int c = 0; c < P.nX; ++c) issue_l0(wp++, c);
            while (wp & (kF2WSlots - 1)) wskipp(wp++);
        }
        for (int tile_iter = 0; tile_of(tile_iter) >= 0; ++tile_iter) {
            const uint32_t tp = (uint32_t)(tile_iter & 1);
            const bool has_next = tile_of(tile_iter + 1) >= 0;
            int c_next = 0;
            // ---- chain of this tile, layer-0 chunks of the next tile behind the stages
            for (int s = 1; s <= sc_n2; ++s) {
                const F2Stage& st = P.st[s];
                const long p_ = wp++;
                const uint32_t wb = wgetp(p_);
                mbar_wait(&tready[s - 1], tp);
                tc_fence_after();
                const uint32_t part_b = st.rows * 128;
                const Opnd B = opnd_mn(tile_a(s - 1), P.tile_rows[s - 1] * 128, 8192);
                const int nk = st.kin >> 4;
                if (st.kind == 1) issue_e(tm + T_CH, opnd_mn(wb, part_b, part_b), B, nk, kF2Jets, false);
                else if (st.kind == 2) issue_e(tm + T_CH, opnd_k(wb, part_b), B, nk, kF2Jets, false);
                else if (a.heads_t) {
                    // heads with the JETS on the lanes: D[64 jets x 16] = H_{n-1}^T (A: the activation tile, MN-major, M = jets) x
                    // W (B: the image block, MN-major, N = out).  An M = 64 accumulator puts jet i on lane (i % 16) + 32 (i / 16),
                    // so the Sampling epilogue spreads over all 16 worker warps with the KLD sum thread-local (features on the
                    // lanes it sits on 10 lanes of ONE quarter: 4.2 k cycles per tile).  A forward pass only: the latent backward
                    // reads mu / log_var with the latent index on the lanes.
                    issue_e(tm + T_MU, B, opnd_mn(wb, 2 * part_b, part_b), nk, 16, false, 64);
                    issue_e(tm + T_LV, B, opnd_mn(wb + part_b, 2 * part_b, part_b), nk, 16, false, 64);
                } else {
                    issue_e(tm + T_MU, opnd_mn(wb, 2 * part_b, part_b), B, nk, kF2Jets, false);
                    issue_e(tm + T_LV, opnd_mn(wb + part_b, 2 * part_b, part_b), B, nk, kF2Jets, false);
                }
                wrelp(p_);
                commit_e(&acc[s]);
                for (int i = sc_cnt(s); i > 0; --i, ++c_next) { if (has_next) issue_l0(wp, c_next); else wskipp(wp); ++wp; }
            }
            if (wp & 1) wskipp(wp++);
            // ---- output layer: 128-feature chunks into two alternating accumulators
            {
                const F2Stage& st = P.st[so];
                const uint32_t part_b = st.rows * 128;
                mbar_wait(&tready[sc_n2], tp);
                tc_fence_after();
                for (int c = 0; c < P.nC; ++c) {
                    const uint32_t b = yk & 1, r = yk >> 1;
                    if (r >= 1) { mbar_wait(&yempty[b], (r - 1) & 1); tc_fence_after(); }
                    const long p0 = wp + 2 * c;
                    wgetp(wp + min(2 * c + 1, P.nX - 1));
                    issue_e(tm + T_Y0 + 64 * b, opnd_mn(slot_a(p0), part_b, P.wslot_bytes),
                            opnd_mn(tile_a(sc_n2), P.tile_rows[sc_n2] * 128, 8192), st.kin >> 4, kF2Jets, false);
                    commit_e(&yfull[b]);
                    ++yk;
                    for (int b2 = 0; b2 < 2 && 2 * c + b2 < P.nX; ++b2) wrelp(p0 + b2);
                }
                wp += P.nX;
            }
            while (wp & (kF2WSlots - 1)) wskipp(wp++);
        }
        __syncwarp();
    } else if (warp == kF3Workers + 1) {
        // ======================================================================================= weight producer
        if (lane == 0) {
            const long total = (long)my_tiles * nops;
            for (long g = 0; g < total; ++g) {
                const int slot = (int)(g & (kF2WSlots - 1)), k = (int)(g % nops);
                if (g >= kF2WSlots) mbar_wait(&wempty[slot], (uint32_t)(((g >> 2) - 1) & 1));
                const uint32_t bytes = (!is_train && k >= n_fwd_ops) ? 16u : P.op_bytes[k];
                mbar_expect_tx(&wfull[slot], bytes);
                bulk_g2s(sm + P.wring_off + slot * P.wslot_bytes, a.wimg + P.op_off[k], bytes, &wfull[slot]);
            }
        }
        __syncwarp();
    } else if (warp == kF3Workers) {
        // ======================================================================================= MMA issuer
        // every lane runs this code with identical values (see f3::mma_e)
        {
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base_s, 0);
            const uint32_t smu = __shfl_sync(0xffffffffu, smem_u32(sm), 0);      // shared-window address of the aligned base
            auto tile_a = [&](int k) { return smu + (uint32_t)toff(k); };
            auto xslot_a = [&](uint32_t b) { return smu + (uint32_t)xslot_off((int)b); };
            const uint32_t gy_a = smu + (uint32_t)P.s_off;
            long gbase = 0;                                  // ops consumed by earlier tiles
            uint32_t xk = 0, yk = 0, gk = 0, dk = 0;         // cumulative uses of the X ring, y slots, g_y slot, dW^T accumulators
            int tile_iter = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++tile_iter) {
                const uint32_t tp = (uint32_t)(tile_iter & 1);
                int wseen = 0;
                auto wslot_addr = [&](int k) { return smu + (uint32_t)P.wring_off + (uint32_t)((gbase + k) & (kF2WSlots - 1)) * (uint32_t)P.wslot_bytes; };
                auto wget = [&](int k) -> uint32_t {        // all ops <= k of this tile have landed; returns the slot of op k
                    for (; wseen <= k; ++wseen) mbar_wait(&wfull[(gbase + wseen) & (kF2WSlots - 1)], (uint32_t)(((gbase + wseen) >> 2) & 1));
                    return wslot_addr(k);
                };
                auto wrel = [&](int k) { commit_e(&wempty[(gbase + k) & (kF2WSlots - 1)]); };   // slot free once the MMAs issued so far complete
                auto wskip = [&](int k) { wget(k); arrive_e(&wempty[(gbase + k) & (kF2WSlots - 1)]); };   // op without a consumer
                auto dw_acquire = [&]() -> uint32_t {        // next dW^T accumulator, drained by the workers
                    const uint32_t b = dk % ndw, r = dk / ndw;
                    if (r >= 1) { mbar_wait(&dwempty[b], (r - 1) & 1); tc_fence_after(); }
                    return tm + kF3_TDW + b * dwc;
                };
                auto dw_commit = [&]() { commit_e(&dwfull[dk % ndw]); ++dk; };

                // ---- forward, stage 0
                {
                    const F2Stage& st = P.st[0];
                    for (int c = 0; c < P.nX; ++c, ++xk) {
                        const uint32_t wb = wget(P.op_first[0] + c);
                        const uint32_t xb_ = xk & xs_mask;
                        mbar_wait(&xfull[xb_], (xk >> xs_log) & 1);
                        tc_fence_after();
                        stamp2();
                        const int nk = min(4, (D - 64 * c + 15) >> 4);
                        issue2_e(tm + T_L0A, tm + T_L0B, opnd_k(wb, st.rows * 128), opnd_k(xslot_a(xb_), 8192), nk, kF2Jets, c > 0, c > 0);
                        commit_e(&xempty[xb_]);
                        wrel(P.op_first[0] + c);
                        stamp2();
                    }
                    commit_e(&acc[0]);
                }
                // ---- forward chain
                for (int s = 1; s <= 2 * n; ++s) {
                    const F2Stage& st = P.st[s];
                    const uint32_t wb = wget(P.op_first[s]);
                    mbar_wait(&tready[s - 1], tp);
                    tc_fence_after();
                    stamp2();
                    const uint32_t part_b = st.rows * 128;
                    const Opnd B = opnd_mn(tile_a((s - 1)), P.tile_rows[s - 1] * 128, 8192);
                    const int nk = st.kin >> 4;
                    if (st.kind == 1) issue_e(tm + T_CH, opnd_mn(wb, part_b, part_b), B, nk, kF2Jets, false);
                    else if (st.kind == 2) issue_e(tm + T_CH, opnd_k(wb, part_b), B, nk, kF2Jets, false);
                    else if (SCORE && a.heads_t) {
                        // heads with the JETS on the lanes: D[64 jets x 16] = H_{n-1}^T (A: the activation tile, MN-major, M = jets) x
                        // W (B: the image block, MN-major, N = out).  An M = 64 accumulator puts jet i on lane (i % 16) + 32 (i / 16),
                        // so the Sampling epilogue runs on four warps of four different schedulers with the KLD sum thread-local
                        // (features on the lanes it sits on 10 lanes of ONE quarter: 4.2 k cycles per tile).  A forward pass only:
                        // the latent backward reads mu / log_var with the latent index on the lanes.
                        issue_e(tm + T_MU, B, opnd_mn(wb, 2 * part_b, part_b), nk, 16, false, 64);
                        issue_e(tm + T_LV, B, opnd_mn(wb + part_b, 2 * part_b, part_b), nk, 16, false, 64);
                    } else {                                                 // heads: mean block, log_var block
                        issue_e(tm + T_MU, opnd_mn(wb, 2 * part_b, part_b), B, nk, kF2Jets, false);
                        issue_e(tm + T_LV, opnd_mn(wb + part_b, 2 * part_b, part_b), B, nk, kF2Jets, false);
                    }
                    wrel(P.op_first[s]);
                    commit_e(&acc[s]);
                    stamp2();
                }
                // ---- output layer (+ its backward products in training)
                {
                    const F2Stage& st = P.st[so];
                    const uint32_t part_b = st.rows * 128;
                    const int ob = P.op_first[so];
                    if (ob - 1 != P.op_first[2 * n]) wskip(ob - 1);                           // alignment dummy
                    mbar_wait(&tready[2 * n], tp);
                    tc_fence_after();
                    auto issue_y = [&](int c) {
                        const uint32_t b = yk & 1, r = yk >> 1;
                        if (r >= 1) { mbar_wait(&yempty[b], (r - 1) & 1); tc_fence_after(); }
                        wget(ob + min(2 * c + 1, P.nX - 1));
                        issue_e(tm + T_Y0 + 64 * b, opnd_mn(wslot_addr(ob + 2 * c), part_b, P.wslot_bytes),
                              opnd_mn(tile_a((2 * n)), P.tile_rows[2 * n] * 128, 8192), st.kin >> 4, kF2Jets, false);
                        commit_e(&yfull[b]);
                        ++yk;
                        // forward pass: nothing else reads the chunk's weight ops, so they leave the ring with THIS product (released
                        // after issue_y(c + 1) they waited for the next product and the ring ran one chunk late)
                        if (!is_train) for (int b2 = 0; b2 < 2 && 2 * c + b2 < P.nX; ++b2) wrel(ob + 2 * c + b2);
                    };
                    stamp2();
                    issue_y(0);
                    stamp2();
                    for (int c = 0; c < P.nC; ++c) {
                        if (c + 1 < P.nC) issue_y(c + 1);
                        stamp2();
                        if (is_train) {
                            mbar_wait(&gyfull[c], tp);
                            tc_fence_after();
                            stamp2();
                            // g_dn += W_out[:, chunk] x g_y[chunk]   (A: K-major rows = in features; B: MN-major, K = chunk features)
                            for (int b = 0; b < 2 && 2 * c + b < P.nX; ++b) {
                                const int nk = min(4, (D - 64 * (2 * c + b) + 15) >> 4);
                                issue_e(tm + T_CH, opnd_k(wslot_addr(ob + 2 * c + b), part_b), opnd_mn(gy_a + b * 8192, 16384, 8192), nk, kF2Jets, c > 0 || b > 0);
                            }
                            // the chunk's weight ops are free once these MMAs complete: the ring can bring in chunk c+2
                            for (int b = 0; b < 2 && 2 * c + b < P.nX; ++b) wrel(ob + 2 * c + b);
                            // dW_out^T[chunk] = g_y (A: K-major, K = jets) x d_n (B: K-major, N = in features)
                            const uint32_t tdw = dw_acquire();
                            issue_e(tdw, opnd_k(gy_a, 16384), opnd_k(tile_a((2 * n)), P.tile_rows[2 * n] * 128), 4, st.kin, false);
                            commit_e(&gyempty); ++gk;
                            dw_commit();
                            stamp2();
                        }
                    }
                    if (is_train) commit_e(&accB[so]);
                }
                if (is_train) {
                    // ---- backward chain: dW_s^T first (it reads H_{s-1}, which the workers overwrite with G_{s-1} next)
                    for (int s = 2 * n; s >= 1; --s) {
                        const F2Stage& st = P.st[s];
                        const uint32_t wb = wget(P.op_bwd[s]);
                        mbar_wait(&gready[s], tp);
                        tc_fence_after();
                        const uint32_t part_b = st.rows * 128;
                        const uint32_t g_hi = tile_a((s)), g_lo = P.tile_rows[s] * 128;
                        const Opnd G = opnd_mn(g_hi, g_lo, 8192);
                        // G_{s-1}^T = W_s x G_s^T
                        if (st.kind == 1) issue_e(tm + T_CH, opnd_k(wb, part_b), G, st.kout >> 4, kF2Jets, false);
                        else if (st.kind == 2) issue_e(tm + T_CH, opnd_mn(wb, part_b, part_b), G, st.kout >> 4, kF2Jets, false);
                        else {
                            issue_e(tm + T_CH, opnd_k(wb, 2 * part_b), G, 1, kF2Jets, false);                                      // W_mu g_mu
                            issue_e(tm + T_CH, opnd_k(wb + part_b, 2 * part_b), opnd_mn(g_hi + 2048, g_lo, 8192), 1, kF2Jets, true);   // + W_lv g_lv
                        }
                        // dW_s^T = G_s (A: K-major, K = jets) x H_{s-1} (B: K-major, N = in features)
                        const uint32_t tdw = dw_acquire();
                        issue_e(tdw, opnd_k(g_hi, g_lo), opnd_k(tile_a((s - 1)), P.tile_rows[s - 1] * 128), 4, st.kin, false);
                        wrel(P.op_bwd[s]);
                        commit_e(&accB[s]);
                        dw_commit();
                    }
                    // ---- stage 0: dW_0^T = G_0 x X, one 64-feature chunk of re-staged X per dW^T accumulator
                    mbar_wait(&gready[0], tp);
                    tc_fence_after();
                    for (int c = 0; c < P.nX; ++c, ++xk) {
                        mbar_wait(&xfull[xk & 1], (xk >> 1) & 1);
                        tc_fence_after();
                        const uint32_t tdw = dw_acquire();
                        issue_e(tdw, opnd_k(tile_a((0)), P.tile_rows[0] * 128), opnd_mn(xslot_a(xk & 1), 8192, 8192), 4, 64, false);
                        commit_e(&xempty[xk & 1]);
                        dw_commit();
                    }
                }
                for (int k = (is_train ? P.op_bwd[1] : n_fwd_ops - 1) + 1; k < nops; ++k) wskip(k);   // trailing dummies / shrunk ops
                gbase += nops;
            }
        }
        __syncwarp();
    } else if (warp >= kF3Workers + 2) {
        if constexpr (SCORE) {
            // =================================================================================== X stager warps (SCORE)
            // X chunk c of a tile = [64 jets x features 64c..64c+63] fp32 -> bf16 hi/lo, K-major rows = jets, into the 2-slot X ring.
            // Nothing here depends on the workers: the ring is freed by the layer-0 MMAs, so the stagers (and the issuer behind them)
            // run layer 0 of tile t+1 under the output epilogues of tile t.
            const int st_id = tid - (kF3Workers + 2) * 32;                  // 0 .. 32 kF3Stagers - 1
            uint32_t xk = 0;
            int n_dbg3 = 0;                                                  // stager stamps (CTA 0, lane 0 of the first stager warp): a.dbg[832] = count, [833..899]
            auto stamp3 = [&](int it_) { if (a.dbg && blockIdx.x == 0 && st_id == 0 && it_ >= dbg_skip && n_dbg3 < 66) a.dbg[833 + n_dbg3++] = clock64(); };
            if (st_id == 0 && a.tile_ctr) { tile_pub(0); tile_pub(1); }
            for (int it = 0; ; ++it) {
                const int tile = tile_of(it);
                if (tile < 0) break;
                if (st_id == 0 && a.tile_ctr && tile_of(it + 1) >= 0) tile_pub(it + 2);     // the queue ends at its first -1: nobody asks beyond it
                int local0, nvalid;
                const float* xa = tile_src(tile, local0, nvalid);
                {   // pull the rows of the tile after this one towards L2
                    const int nt = tile_of(it + 1);
                    int l0, nv;
                    const char* nx = reinterpret_cast<const char*>(tile_src(max(nt, 0), l0, nv));
                    const long nbytes = nt >= 0 ? (long)nv * D * 4 : 0;
                    for (long o = (long)st_id * 128; o < nbytes; o += kF3Stagers * 32 * 128) prefetch_l2(nx + o);
                }
                // a chunk is two halves of 512 float4; the loads of the next half are in flight while this one is converted
                constexpr int kSt = kF3Stagers * 32, kIt = (512 + kSt - 1) / kSt;      // float4 per thread and half
                auto ldh = [&](int hc, float4 (&r)[kIt]) {               // hc = 2 * chunk + half
                    const int c = hc >> 1;
#pragma unroll
                    for (int it = 0; it < kIt; ++it) {
                        const int ih = st_id + kSt * it, idx = ih + 512 * (hc & 1), rr = idx >> 4, c4 = (idx & 15) << 2, col = 64 * c + c4;
                        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (c < P.nX && rr < nvalid && ih < 512) {
                            const float* p = xa + (long)rr * D + col;
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
                auto sth = [&](int hc, const float4 (&r)[kIt]) {
                    const uint32_t b = xk & xs_mask, rnd = xk >> xs_log;
                    if (!(hc & 1) && rnd >= 1) mbar_wait(&xempty[b], (rnd - 1) & 1);
                    unsigned char* hi = xslot((int)b); unsigned char* lo = hi + 8192;
#pragma unroll
                    for (int it = 0; it < kIt; ++it) {
                        const int ih = st_id + kSt * it, idx = ih + 512 * (hc & 1), rr = idx >> 4, c4 = (idx & 15) << 2;
                        if (ih < 512) {
                            uint32_t ha, la, hb, lb;
                            split2(r[it].x, r[it].y, ha, la); split2(r[it].z, r[it].w, hb, lb);
                            const uint32_t off = sw128_off(rr, c4);
                            *reinterpret_cast<uint2*>(hi + off) = make_uint2(ha, hb);
                            *reinterpret_cast<uint2*>(lo + off) = make_uint2(la, lb);
                        }
                    }
                    if (hc & 1) { fence_async_smem(); warp_arrive(&xfull[b], lane); ++xk; }
                };
                // Fast versions for whole tiles whose rows are 16-byte aligned (D % 4 == 0): the slot geometry of a thread is
                // k-invariant (row r0 + 4 it + 32 half, column c4), so a load is one add + LDG.128 and a store one add + two STS.64.
                // The stagers share their schedulers with the worker warps, whose output epilogues are issue bound: at ~40 instructions
                // per float4 they needed 3.7-7 k cycles per chunk and paced the whole pass (bench/score_stamps.py).
                const bool fast_tile = kSt == 64 && nvalid == kF2Jets && (D & 3) == 0 && (reinterpret_cast<uintptr_t>(xa) & 15) == 0;
                const int r0 = st_id >> 4, c4f = (st_id & 15) << 2;
                const float* rowbase = xa + (long)r0 * D + c4f;
                const long d4 = 4L * D, d32 = 32L * D;
                const uint32_t sw0 = (uint32_t)r0 * 128u + (uint32_t)(((((c4f >> 3) ^ r0) & 7) << 4) | ((c4f & 7) << 1));
                const uint32_t sw1 = (uint32_t)r0 * 128u + (uint32_t)(((((c4f >> 3) ^ (r0 | 4)) & 7) << 4) | ((c4f & 7) << 1));
                auto ldf = [&](int hc, float4 (&r)[kIt]) {
                    const float* p = rowbase + (hc & 1) * d32 + 64 * (hc >> 1);
                    const bool on = 64 * (hc >> 1) + c4f < D;                 // last chunk: columns beyond D are zero (D % 4 == 0 here)
#pragma unroll
                    for (int it = 0; it < kIt; ++it) r[it] = on ? __ldg(reinterpret_cast<const float4*>(p + it * d4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                };
                auto stf = [&](int hc, const float4 (&r)[kIt]) {
                    const uint32_t b = xk & xs_mask, rnd = xk >> xs_log;
                    if (!(hc & 1) && rnd >= 1) mbar_wait(&xempty[b], (rnd - 1) & 1);
                    unsigned char* hi = xslot((int)b) + (hc & 1) * 4096; unsigned char* lo = hi + 8192;
#pragma unroll
                    for (int it = 0; it < kIt; ++it) {
                        uint32_t ha, la, hb, lb;
                        split2(r[it].x, r[it].y, ha, la); split2(r[it].z, r[it].w, hb, lb);
                        const uint32_t off = (uint32_t)it * 512u + ((it & 1) ? sw1 : sw0);
                        *reinterpret_cast<uint2*>(hi + off) = make_uint2(ha, hb);
                        *reinterpret_cast<uint2*>(lo + off) = make_uint2(la, lb);
                    }
                    if (hc & 1) { fence_async_smem(); warp_arrive(&xfull[b], lane); ++xk; }
                };
                auto chunk_fast = [&](int hc) { return fast_tile && (hc >> 1) < P.nX; };
                auto ld = [&](int hc, float4 (&r)[kIt]) { if (chunk_fast(hc)) ldf(hc, r); else ldh(hc, r); };
                auto st = [&](int hc, const float4 (&r)[kIt]) { if (chunk_fast(hc)) stf(hc, r); else sth(hc, r); };
                float4 ra[kIt], rb[kIt];
                stamp3(it);
                ld(0, ra);
                for (int hc = 0; hc < 2 * P.nX; hc += 2) {
                    ld(hc + 1, rb); st(hc, ra);
                    ld(hc + 2, ra); st(hc + 1, rb);
                    stamp3(it);
                }
            }
            if (a.dbg && blockIdx.x == 0 && st_id == 0) a.dbg[832] = n_dbg3;
        }
        // ======================================================================================= drain warps
        // dW^T accumulators (lane = output feature, column = input feature) -> rows of the Keras [in, out] kernel gradients
        // in this CTA's partial-gradient slice; a warp writes 32 consecutive outputs per row (one 128-byte segment per
        // store).  Order = the issuer's: output chunks, chain stages 2n..1, layer-0 chunks.
        if (is_train) {
            uint32_t dk = 0;
            int tile_iter = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++tile_iter) {
                const uint32_t tp = (uint32_t)(tile_iter & 1);
                float* part = a.parts + (long)blockIdx.x * P.Ppad;
                auto put_dw = [&](float* dw, long ld, int i0, int in, const float* v) {
                    float* d = dw + (long)i0 * ld;
                    const int cnt = in - i0;                             // warp-uniform
                    if (tile_iter == 0) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) { if (j < cnt) *d = v[j]; d += ld; }
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) { if (j < cnt) atomicAdd(d, v[j]); d += ld; }   // RED: no load -> add -> store chain
                    }
                };
                auto put_db = [&](float* d, int buf, int row) {
                    const float t = (db_at(buf, 0, row) + db_at(buf, 1, row)) + (db_at(buf, 2, row) + db_at(buf, 3, row));
                    if (tile_iter) atomicAdd(d, t); else *d = t;
                };
                auto dw_wait = [&]() -> uint32_t {
                    const uint32_t b = dk % ndw, r = dk / ndw;
                    mbar_wait(&dwfull[b], r & 1);
                    tc_fence_after();
                    return t_lane + kF3_TDW + b * dwc;
                };
                auto dw_release = [&]() { tc_fence_before(); warp_arrive(&dwempty[dk % ndw], lane); ++dk; };
                // ---- output layer: chunk c = features 128c.. (bias-gradient parts of loss(c) are behind gyfull[c])
                {
                    const F2Stage& st = P.st[so];
                    for (int c = 0; c < P.nC; ++c) {
                        mbar_wait(&gyfull[c], tp);
                        const uint32_t tdw = dw_wait();
                        const int f = 128 * c + frow;
                        const bool valid = f < D;
                        for (int g = 0; g < (st.kin >> 4); ++g) {
                            float v[16];
                            tmem_ld16(tdw + 16 * g, v);
                            if (valid) put_dw(part + st.seg_w[0] + f, D, 16 * g, st.in, v);
                        }
                        if (valid) put_db(part + st.seg_b[0] + f, 3 + c % 3, frow);
                        dw_release();
                    }
                }
                // this CTA's share of the output layer's gradient is final: the update kernel's blocks for those parameters may start
                // (release: every lane's stores / reductions are performed at the L2 before lane 0 counts the warp in)
                const bool last_tile = tile + (int)gridDim.x >= a.n_tiles;
                if (a.grp_ctr && last_tile) { __threadfence(); __syncwarp(); if (lane == 0) atomicAdd(a.grp_ctr, 1u); }
                // ---- chain stages (bias-gradient parts of G_s are behind gready[s])
                for (int s = 2 * n; s >= 1; --s) {
                    const F2Stage& st = P.st[s];
                    mbar_wait(&gready[s], tp);
                    const uint32_t tdw = dw_wait();
                    const int top = st.seg_lane0[st.nseg - 1] + st.seg_n[st.nseg - 1];
                    if (32 * q < top) {
                        int sg = -1;
#pragma unroll
                        for (int k = 0; k < 2; ++k) if (k < st.nseg && frow >= st.seg_lane0[k] && frow < st.seg_lane0[k] + st.seg_n[k]) sg = k;
                        const int o = sg >= 0 ? frow - st.seg_lane0[sg] : 0, ld = sg >= 0 ? st.seg_n[sg] : 0;
                        float* dw = part + (sg >= 0 ? st.seg_w[sg] : 0) + o;
                        for (int g = 0; g < (st.kin >> 4); ++g) {
                            float v[16];
                            tmem_ld16(tdw + 16 * g, v);
                            if (sg >= 0) put_dw(dw, ld, 16 * g, st.in, v);
                        }
                        if (sg >= 0) put_db(part + st.seg_b[sg] + o, s % 3, frow);
                    }
                    dw_release();
                }
                if (a.grp_ctr && last_tile) { __threadfence(); __syncwarp(); if (lane == 0) atomicAdd(a.grp_ctr + 1, 1u); }
                // ---- layer 0: one 64-feature chunk of dW_0^T per accumulator.  Rows 64c..64c+63 of the Keras [D, out_0]
                // kernel gradient are one contiguous block, so each chunk is transposed through shared memory (tiles 1..2n are
                // dead) and leaves as ONE bulk copy (bulk reduce-add on later tiles) instead of 16 scalar stores per thread:
                // an SM stores ~17 B/clk through the LSU, which is what bounded this phase.
                if (P.stg0_off >= 0) {
                    const F2Stage& st = P.st[0];
                    const bool leader = warp == kF3Workers + 2 && lane == 0;
                    const int out0 = st.out;
                    mbar_wait(&gready[0], tp);
                    for (int c = 0; c < P.nX; ++c) {
                        const uint32_t tdw = dw_wait();
                        float* stg = reinterpret_cast<float*>(sm + P.stg0_off) + (c & 1) * 64 * out0;
                        if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");     // chunk c-2 has left this buffer
                        asm volatile("bar.sync 2, %0;" ::"n"(kF3Drains * 32) : "memory");
                        if (32 * q < out0) {
                            const bool valid = frow < out0;
                            for (int g = 0; g < 4; ++g) {
                                float v[16];
                                tmem_ld16(tdw + 16 * g, v);
                                if (valid) {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) stg[(16 * g + j) * out0 + frow] = v[j];
                                }
                            }
                            if (valid && c == 0) put_db(part + st.seg_b[0] + frow, 0, frow);
                        }
                        fence_async_smem();
                        dw_release();
                        asm volatile("bar.sync 2, %0;" ::"n"(kF3Drains * 32) : "memory");
                        if (leader) {
                            const uint32_t bytes = (uint32_t)(min(64, D - 64 * c) * out0 * 4);
                            float* dst = part + st.seg_w[0] + (long)64 * c * out0;
                            if (tile_iter == 0)
                                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(stg)), "r"(bytes) : "memory");
                            else
                                asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(stg)), "r"(bytes) : "memory");
                            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                        }
                    }
                    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");              // stores performed before the CTA moves on
                    asm volatile("bar.sync 2, %0;" ::"n"(kF3Drains * 32) : "memory");
                } else {
                    const F2Stage& st = P.st[0];
                    mbar_wait(&gready[0], tp);
                    for (int c = 0; c < P.nX; ++c) {
                        const uint32_t tdw = dw_wait();
                        if (32 * q < st.out) {
                            const bool valid = frow < st.out;
                            for (int g = 0; g < 4; ++g) {
                                float v[16];
                                tmem_ld16(tdw + 16 * g, v);
                                if (valid) put_dw(part + st.seg_w[0] + frow, st.out, 64 * c + 16 * g, D, v);
                            }
                            if (valid && c == 0) put_db(part + st.seg_b[0] + frow, 0, frow);
                        }
                        dw_release();
                    }
                }
            }
        }
    } else {
        // ======================================================================================= workers
        uint32_t xk = 0, yk = 0, gk = 0;
        int tile_iter = 0;
        for (;; ++tile_iter) {
            const int tile = tile_of(tile_iter);
            if (tile < 0) break;
            const uint32_t tp = (uint32_t)(tile_iter & 1);
            dbg_on = tile_iter >= dbg_skip;
            int local0, nvalid;                                       // local0: row inside its input = the row the noise is drawn for
            const float* xa = tile_src(tile, local0, nvalid);
            // row in the per-jet arrays ([bkg ; OoD], compact); recomputed where it is used instead of living through the tile
            auto row0_of = [&]() -> long { int l0, nv; tile_src(tile, l0, nv); return (!TWOIN || tile * kF2Jets == l0) ? (long)l0 : a.B1 + l0; };
            // jets 32..63 of a paired tile sit at xb + j D; column j is a real jet iff (j & cmask) < nvalid
            const float* xb = PAIRED ? a.x2 + (long)(local0 - 32) * D : xa;
            auto cidx = [&](int j) { return PAIRED ? (j & 31) : j; };             // index of column j inside its input
            auto colv = [&](int j) { return cidx(j) < nvalid; };
            if (PAIRED) {
                // s_w: coefficient of the reconstruction term (w_bkg | 0); s_wk: beta w_bkg | lambda w_OoD -- the latent backward
                // turns them into the KLD coefficients of the pair once both KLDs are known
                if (tid < 64) {
                    const bool on = (tid & 31) < nvalid;
                    const float wv = on ? (tid < 32 ? a.w[local0 + tid] : a.w2[local0 + tid - 32]) : 0.f;
                    s_w[tid] = tid < 32 ? wv : 0.f; s_wk[tid] = (tid < 32 ? a.beta : a.lamb) * wv;
                }
            } else if (tid < 64) {
                // coefficients of the two loss terms: (w, beta w) of get_losses, unit weights for the raw pre-pass, or the per-jet
                // derivatives of the outlier-exposure loss (combine_kernel) when the step trains on bkg + OoD rows at once
                float cr = 0.f, ck = 0.f;
                if (tid < nvalid && !SCORE) {
                    const long row0 = row0_of();
                    if (TWOIN && a.c_rec) { cr = a.c_rec[row0 + tid]; ck = a.c_kld[row0 + tid]; }
                    else if (TWOIN) { cr = 1.f; ck = 1.f; }
                    else { cr = a.w[row0 + tid]; ck = a.beta * cr; }
                }
                s_w[tid] = cr; s_wk[tid] = ck;
            }
            if constexpr (!SCORE) {   // pull the next tile's rows of X towards L2 (SCORE: the stagers do)
                int l0, nv;
                const char* nx = reinterpret_cast<const char*>(tile_src(tile + (int)gridDim.x, l0, nv));
                const long nbytes = (tile + (int)gridDim.x < a.n_tiles) ? (long)nv * D * 4 : 0;
                for (long o = (long)tid * 128; o < nbytes; o += kF3Workers * 32 * 128) prefetch_l2(nx + o);
                if (PAIRED) {
                    const char* nx2 = reinterpret_cast<const char*>(a.x2 + (long)l0 * D);
                    for (long o = (long)tid * 128; o < nbytes; o += kF3Workers * 32 * 128) prefetch_l2(nx2 + o);
                }
            }
            // reparameterisation noise of the tile -> s_eps[l][jet]: external eps, or one Philox counter (row, quad, step)
            if (a.eps) {
                for (int i = tid; i < 64 * L; i += kF3Workers * 32) {
                    const int j = i / L, l = i - j * L;
                    s_eps[l * 64 + j] = colv(j) ? __ldg(a.eps + (local0 + cidx(j)) * L + l) : 0.f;
                }
            } else if (tid < 64 * ((L + 3) >> 2)) {
                const int j = tid & 63, quad = tid >> 6;
                float nrm[4];
                philox_normal4(a.seed, a.row_offset + (unsigned long long)(local0 + cidx(j)), (unsigned)quad, a.step, nrm);
#pragma unroll
                for (int t = 0; t < 4; ++t) if (4 * quad + t < L) s_eps[(4 * quad + t) * 64 + j] = nrm[t];
            }

            // X chunk c = [64 jets x features 64c..64c+63]: ldx pulls this thread's two float4 into registers, stx writes them
            // to a ring slot as bf16 hi/lo (K-major rows = jets)
            auto ldx = [&](int c, float4 (&r)[2]) {
#pragma unroll
                for (int it = 0; it < 2; ++it) {
                    const int idx = tid + kF3Workers * 32 * it, rr = idx >> 4, c4 = (idx & 15) << 2, col = 64 * c + c4;
                    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (c < P.nX && colv(rr)) {
                        const float* p = ((PAIRED && it) ? xb : xa) + (long)rr * D + col;        // it = 1: rows 32..63
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
                    const int idx = tid + kF3Workers * 32 * it, rr = idx >> 4, c4 = (idx & 15) << 2;
                    uint32_t ha, la, hb, lb;
                    split2(r[it].x, r[it].y, ha, la); split2(r[it].z, r[it].w, hb, lb);
                    const uint32_t off = sw128_off(rr, c4);
                    *reinterpret_cast<uint2*>(hi + off) = make_uint2(ha, hb);
                    *reinterpret_cast<uint2*>(lo + off) = make_uint2(la, lb);
                }
            };
            // stage chunk c into the next ring slot (waits until the MMAs that read the slot's previous content are done)
            auto stage_x = [&](float4 (&r)[2]) {
                const uint32_t b = xk & 1, rr = xk >> 1;
                if (rr >= 1) mbar_wait(&xempty[b], (rr - 1) & 1);
                stx(b, r);
                fence_async_smem();
                warp_arrive(&xfull[b], lane);
                ++xk;
            };
            // smem tile writes + TMEM reads of this thread are done -> one arrival per warp
            auto publish = [&](uint64_t* bar) { fence_async_smem(); tc_fence_before(); warp_arrive(bar, lane); };
            auto wait_acc = [&](uint64_t* bar, uint32_t parity) { mbar_wait(bar, parity); tc_fence_after(); };

            // chain accumulator -> bias + ReLU -> tile s (pad rows zero)
            auto epi_act = [&](int s, float bias, bool two, uint32_t col = T_CH) {
                const F2Stage& st = P.st[s];
                if (32 * q >= P.tile_rows[s]) return;
                float v[16];
                tmem_ld16(t_lane + col + jc0, v);
                if (two) {
                    float u[16];
                    tmem_ld16(t_lane + T_L0B + jc0, u);
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += u[j];
                }
                if (frow < P.tile_rows[s]) {
                    // rows out..tile_rows-1 hold exact zeros (zero image rows, bias_of() = 0 there)
                    if (st.relu) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j] + bias, 0.f);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = v[j] + bias;
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
                    db_at((s - 1) % 3, h, frow) = dbs;
                }
            };
            worker_sync();                                               // s_w / s_eps visible; previous tile's scratch reads done
            // =========================================================== forward, stage 0: X streamed in 64-feature chunks
            {
                if constexpr (!SCORE) {                                   // SCORE: the stager warps fill the X ring
                    float4 ra[2], rb[2];
                    ldx(0, ra); ldx(1, rb);
                    for (int c = 0; c < P.nX; c += 2) {
                        stage_x(ra); ldx(c + 2, ra);
                        if (c + 1 < P.nX) { stage_x(rb); ldx(c + 3, rb); }
                        if (c == 0 && tile_iter == 0) asm volatile("griddepcontrol.wait;" ::: "memory");   // see late_wait
                    }
                }
                const float bias0 = bias_of(0);
                wait_acc(&acc[0], tp);
                stamp();
                epi_act(0, bias0, D > 16, T_L0A);
                publish(&tready[0]);
                stamp();
            }
            // =========================================================== forward chain: stages 1 .. 2n
            for (int s = 1; s <= 2 * n; ++s) {
                const F2Stage& st = P.st[s];
                float b_mu = 0.f, b_lv = 0.f;
                if (s != n) b_mu = bias_of(s);
                else if (!(SCORE && a.heads_t) && (lane & 15) < L) { b_mu = __ldg(a.theta + st.seg_b[0] + (lane & 15)); b_lv = __ldg(a.theta + st.seg_b[1] + (lane & 15)); }
                if (SCORE && a.heads_t && s == n) {                      // the plan has L <= 16
                    // ---- heads, jets on the lanes (see the issuer): warp (q, h) owns jets 16 q .. 16 q + 15 and latent rows 4 h .. 4 h + 3;
                    // the four partial KLD sums of a jet meet in the per-jet write at the end of the tile
                    float bm[4], bl[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int l = 4 * h + i;
                        bm[i] = l < L ? __ldg(a.theta + st.seg_b[0] + l) : 0.f; bl[i] = l < L ? __ldg(a.theta + st.seg_b[1] + l) : 0.f;
                    }
                    wait_acc(&acc[s], tp);
                    stamp();
                    float mu[4], lv[4];
                    tmem_ld4(t_lane + T_MU + 4 * h, mu);
                    tmem_ld4(t_lane + T_LV + 4 * h, lv);
                    const int j = 16 * q + (lane & 15);
                    float ks = 0.f;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int l = 4 * h + i;
                        const float m = mu[i] + bm[i], v = lv[i] + bl[i];
                        const float sg = __expf(v * 0.5f);
                        const float zz = m + clipv(sg) * s_eps[min(l, L - 1) * 64 + j];
                        const float kk = 1.f + v - clipv(sg * sg) - m * m;
                        mu[i] = l < L ? zz : 0.f;
                        ks += l < L ? kk : 0.f;
                    }
                    if (lane < 16) {
#pragma unroll
                        for (int i = 0; i < 4; i += 2) {
                            uint32_t hh, ll;
                            split2(mu[i], mu[i + 1], hh, ll);
                            const uint32_t o0 = sw128_off(4 * h + i, j), o1 = sw128_off(4 * h + i + 1, j);
                            *reinterpret_cast<unsigned short*>(tile_hi(n) + o0) = (unsigned short)(hh & 0xffffu);
                            *reinterpret_cast<unsigned short*>(tile_hi(n) + o1) = (unsigned short)(hh >> 16);
                            *reinterpret_cast<unsigned short*>(tile_lo(n) + o0) = (unsigned short)(ll & 0xffffu);
                            *reinterpret_cast<unsigned short*>(tile_lo(n) + o1) = (unsigned short)(ll >> 16);
                        }
                        s_db[1024 + 64 * h + j] = ks;                           // s_db[0, 768) holds the metric partials, the rest is free in a forward pass
                    }
                    publish(&tready[s]);
                    stamp();
                    continue;
                }
                wait_acc(&acc[s], tp);
                stamp();
                if (s != n) {
                    epi_act(s, b_mu, false);
                    stamp();
                } else if (q == 0 && P.heads_rep) {
                    // ---- heads: Sampling (models.py:9-15) + KLD (models.py:81-85); z -> rows 0..15 of the latent tile.
                    // Accumulator rows 16 + l replicate rows l (second copy of the latent columns in the weight image), so lane l takes
                    // jets 0..7 of the warp's column group and lane 16 + l jets 8..15: half the serial work of the 10-lane version.
                    const int l = lane & 15, j0 = jc0 + 8 * (lane >> 4);
                    float mu[8], lv[8], kl[8];
                    {   // tcgen05.ld takes one warp-uniform address: every lane reads the 16 columns and keeps its half
                        float m16[16], v16[16];
                        tmem_ld16(t_lane + T_MU + jc0, m16);
                        tmem_ld16(t_lane + T_LV + jc0, v16);
                        const bool hi8 = lane >= 16;
#pragma unroll
                        for (int j = 0; j < 8; ++j) { mu[j] = hi8 ? m16[8 + j] : m16[j]; lv[j] = hi8 ? v16[8 + j] : v16[j]; }
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float zz = 0.f, kk = 0.f;
                        if (l < L) {
                            const float m = mu[j] + b_mu, v = lv[j] + b_lv;
                            const float sg = __expf(v * 0.5f);                      // sigma; exp(v) = sigma^2 (SFU ex2, rel err ~1e-6)
                            zz = m + clipv(sg) * s_eps[l * 64 + j0 + j];
                            kk = 1.f + v - clipv(sg * sg) - m * m;
                        }
                        mu[j] = zz; kl[j] = kk;
                    }
                    store8(tile_hi(n), tile_lo(n), l, j0, mu);
                    const float ks = transpose_reduce8(kl, lane);
                    if ((lane & 8) == 0) s_kld[j0 + (lane & 7)] = -(ks / (float)L) / 2.f;
                } else if (q == 0) {
                    // ---- the same with the latent index on lanes 0..L-1 only (no replicated rows in this plan)
                    float mu[16], lv[16], kl[16];
                    tmem_ld16(t_lane + T_MU + jc0, mu);
                    tmem_ld16(t_lane + T_LV + jc0, lv);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float zz = 0.f, kk = 0.f;
                        if (lane < L) {
                            const float m = mu[j] + b_mu, v = lv[j] + b_lv;
                            const float sg = __expf(v * 0.5f);                      // sigma; exp(v) = sigma^2 (SFU ex2, rel err ~1e-6)
                            zz = m + clipv(sg) * s_eps[lane * 64 + jc0 + j];
                            kk = 1.f + v - clipv(sg * sg) - m * m;
                        }
                        mu[j] = zz; kl[j] = kk;
                    }
                    if (lane < 16) store16(tile_hi(n), tile_lo(n), lane, jc0, mu);
                    const float ks = transpose_reduce16(kl, lane);
                    if (lane < 16) s_kld[jc0 + lane] = -(ks / (float)L) / 2.f;
                }
                publish(&tready[s]);
                stamp();
            }
            if (!SCORE && PAIRED) worker_sync();                       // both KLDs of every pair are in s_kld
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
                const float inv_d = 1.f / (float)D;
                float wsc[SCORE ? 1 : 16];                               // jet weight / D of this thread's 16 jets
                if constexpr (!SCORE) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) wsc[j] = s_w[jc0 + j] * inv_d;
                }
                // column jc0 + j of the tile is a real jet iff j < nj; xcol: x of this thread's first jet at feature frow
                const int nj = max(0, min(16, nvalid - cidx(jc0)));
                const float* xcol = ((PAIRED && jc0 >= 32) ? xb : xa) + (long)jc0 * D + frow;
                const bool want_kj = SCORE && (a.s_kld || a.s_jsd);
                for (int c = 0; c < P.nC; ++c) {
                    const int f = 128 * c + frow;
                    const bool valid = f < D;
                    // this thread's 16 x values of the chunk (L2-resident since stage 0), issued before the accumulator wait
                    float xv[16];
                    {
                        const float* px = xcol + 128 * c;
                        if (valid && nj == 16) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) xv[j] = __ldg(px + j * D);
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) xv[j] = (valid && j < nj) ? __ldg(px + j * D) : 0.f;
                        }
                    }
                    const float bias = valid ? __ldg(a.theta + st.seg_b[0] + f) : 0.f;
                    const uint32_t yb = yk & 1;
                    stamp();
                    wait_acc(&yfull[yb], (yk >> 1) & 1);
                    stamp();
                    float y[16];
                    tmem_ld16(t_lane + T_Y0 + 64 * yb + jc0, y);
                    tc_fence_before();
                    warp_arrive(&yempty[yb], lane);
                    ++yk;
                    stamp();
                    if constexpr (SCORE) {
                        // loss_function (utils.py:346-375) terms of this thread's feature for its 16 jets; Q = clip_values(X_hat)
                        // Columns j >= nj accumulate values that are never written.  Pass 1 forms y + bias and four cheap range
                        // checks over the thread's 16 elements; when they hold (|y|, |x| < 1e6 = the clip bound, |y| and |x + y| > 1e-30 -- anything a
                        // trained model produces) clip_values is the identity, every logarithm below is finite or belongs to a
                        // term that nan_to_num zeroes, and pass 2 is ~20 instructions per element.  Otherwise the thread takes the
                        // exact path (clip_values + the quotient form with its NaN / infinity rules) for this chunk.
                        if (valid) {
                            float ssx = 0.f, ssp = 0.f, mnx = 3.0e38f, mns = 3.0e38f;
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                y[j] += bias;
                                ssx = fmaf(y[j], y[j], ssx); ssp = fmaf(xv[j], xv[j], ssp);
                                mnx = fminf(mnx, fabsf(y[j])); mns = fminf(mns, fabsf(xv[j] + y[j]));
                            }
                            const bool tame = ssx < 1e12f && ssp < 1e12f;     // every |y| and |x| below 1e6: clip_values is the identity
                            if (tame && (!want_kj || (mnx > 1e-30f && mns > 4e-30f))) {
                                if (want_kj) {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) {
                                        const float p = xv[j], qv = y[j], d = p - qv;
                                        if constexpr (!NOMSE) rec_acc[j] = fmaf(d, d, rec_acc[j]);
                                        sc_mae[j] += fabsf(d);
                                        score_kj(p, qv, sc_kl[j], sc_js[j]);
                                    }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) {
                                        const float d = xv[j] - y[j];
                                        if constexpr (!NOMSE) rec_acc[j] = fmaf(d, d, rec_acc[j]);
                                        sc_mae[j] += fabsf(d);
                                    }
                                }
                            } else {
                                if (a.dbg) {   // diagnostics: thread-chunks on the exact path, by failed condition
                                    atomicAdd(reinterpret_cast<unsigned long long*>(&a.dbg[950]), 1ull);
                                    if (!(ssx < 1e12f)) atomicAdd(reinterpret_cast<unsigned long long*>(&a.dbg[951]), 1ull);
                                    if (!(ssp < 1e12f)) atomicAdd(reinterpret_cast<unsigned long long*>(&a.dbg[952]), 1ull);
                                    if (!(mnx > 1e-30f)) atomicAdd(reinterpret_cast<unsigned long long*>(&a.dbg[953]), 1ull);
                                    if (!(mns > 4e-30f)) atomicAdd(reinterpret_cast<unsigned long long*>(&a.dbg[954]), 1ull);
                                }
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    const float p = xv[j], qv = clipv(y[j]), d = p - qv;
                                    y[j] = qv;
                                    if constexpr (!NOMSE) rec_acc[j] = fmaf(d, d, rec_acc[j]);
                                    sc_mae[j] += fabsf(d);
                                    if (want_kj) score_kj_slow(p, qv, sc_kl[j], sc_js[j]);
                                }
                            }
                            if (a.s_xhat) {
                                float* px = a.s_xhat + (row0_of() + jc0) * D + f;
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < nj) px[j * D] = y[j];
                            }
                        }
                    } else {
                        // reconstruction loss (models.py:66,72-78) and its gradient through clip_values: X_hat = clip_values(y)
                        // is 0 for a non-finite y and +-1e6 beyond the clip; the gradient passes only where |y| <= 1e6
                        float dbs = 0.f;
                        if (a.mse) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const float yr = y[j] + bias, ay = fabsf(yr);
                                const float cl = ay < __int_as_float(0x7f800000) ? fminf(fmaxf(yr, -kClip), kClip) : 0.f;
                                const float d = valid ? cl - xv[j] : 0.f;
                                rec_acc[j] = fmaf(d, d, rec_acc[j]);
                                const float gy = ay <= kClip ? 2.f * wsc[j] * d : 0.f;
                                dbs += gy; y[j] = gy;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const float yr = y[j] + bias, ay = fabsf(yr);
                                const float cl = ay < __int_as_float(0x7f800000) ? fminf(fmaxf(yr, -kClip), kClip) : 0.f;
                                const float d = valid ? cl - xv[j] : 0.f;
                                rec_acc[j] += fabsf(d);
                                const float sg = d == 0.f ? 0.f : (d < 0.f ? -wsc[j] : wsc[j]);   // coefficient * sign(d) / D (the OE coefficient can be negative)
                                const float gy = ay <= kClip ? sg : 0.f;
                                dbs += gy; y[j] = gy;
                            }
                        }
                        stamp();
                        if (is_train) {
                            if (gk >= 1) mbar_wait(&gyempty, (gk - 1) & 1);          // MMAs of the previous chunk have read g_y
                            stamp();
                            store16(gy_hi, gy_lo, frow, jc0, y);
                            db_at(3 + c % 3, h, frow) = dbs;
                            fence_async_smem();
                            warp_arrive(&gyfull[c], lane);
                            ++gk;
                        }
                    }
                    stamp();
                }
                if constexpr (!NOMSE) {
                    const float rs = transpose_reduce16(rec_acc, lane);
                    if (lane < 16) s_rec[q * 64 + jc0 + lane] = rs;
                }
                if constexpr (SCORE) {                                       // s_db is free outside training: [3][4 quarters][64 jets]
                    const float r1 = transpose_reduce16(sc_mae, lane), r2 = transpose_reduce16(sc_kl, lane), r3 = transpose_reduce16(sc_js, lane);
                    if (lane < 16) { s_db[q * 64 + jc0 + lane] = r1; s_db[256 + q * 64 + jc0 + lane] = r2; s_db[512 + q * 64 + jc0 + lane] = r3; }
                }
                if (is_train) {
                    wait_acc(&accB[so], tp);
                    epi_grad(so);
                    publish(&gready[2 * n]);
                }
                stamp();
            }
            // =========================================================== backward chain: stages 2n .. 1
            if (is_train) {
                for (int s = 2 * n; s >= 1; --s) {
                    float b_mu = 0.f, b_lv = 0.f;
                    if (s == n + 1 && (lane & 15) < L) { b_mu = __ldg(a.theta + P.st[n].seg_b[0] + (lane & 15)); b_lv = __ldg(a.theta + P.st[n].seg_b[1] + (lane & 15)); }
                    wait_acc(&accB[s], tp);
                    if (s != n + 1) {
                        epi_grad(s);
                    } else if (q == 0 && P.heads_rep) {
                        // ---- latent backward: g_z -> g_mu (rows 0..15), g_logvar (rows 16..31) of the latent tile;
                        // lanes l / 16 + l split the 16 jets of the column group (replicated accumulator rows, see the heads)
                        const float inv_l = 1.f / (float)L;
                        const int l = lane & 15, j0 = jc0 + 8 * (lane >> 4);
                        float dbm = 0.f, dbl = 0.f;
                        float gz[8], mu[8], lv[8];
                        {   // one warp-uniform TMEM address per load: read the 16 columns, keep this lane's half
                            float t16[16];
                            const bool hi8 = lane >= 16;
                            tmem_ld16(t_lane + T_CH + jc0, t16);
#pragma unroll
                            for (int j = 0; j < 8; ++j) gz[j] = hi8 ? t16[8 + j] : t16[j];
                            tmem_ld16(t_lane + T_MU + jc0, t16);
#pragma unroll
                            for (int j = 0; j < 8; ++j) mu[j] = hi8 ? t16[8 + j] : t16[j];
                            tmem_ld16(t_lane + T_LV + jc0, t16);
#pragma unroll
                            for (int j = 0; j < 8; ++j) lv[j] = hi8 ? t16[8 + j] : t16[j];
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float gm = 0.f, gl = 0.f;
                            if (l < L) {
                                const float m = mu[j] + b_mu, v = lv[j] + b_lv;
                                float gkw = s_wk[j0 + j];
                                if (PAIRED) {
                                    const int pc = (j0 + j) & 31;
                                    const bool act = s_kld[pc] - s_kld[pc + 32] + a.margin > 0.f;
                                    gkw = j0 < 32 ? gkw + (act ? s_wk[pc + 32] : 0.f) : (act ? -gkw : 0.f);
                                }
                                const float s_raw = __expf(v * 0.5f), e_raw = s_raw * s_raw;
                                gm = gz[j] + gkw * m * inv_l;
                                gl = gkw * (-(1.f - (clip_pass(e_raw) ? e_raw : 0.f)) * 0.5f * inv_l)
                                   + gz[j] * s_eps[l * 64 + j0 + j] * (clip_pass(s_raw) ? s_raw * 0.5f : 0.f);
                            }
                            mu[j] = gm; lv[j] = gl; dbm += gm; dbl += gl;
                        }
                        dbm += __shfl_xor_sync(0xffffffffu, dbm, 16); dbl += __shfl_xor_sync(0xffffffffu, dbl, 16);
                        store8(tile_hi(n), tile_lo(n), l, j0, mu);
                        store8(tile_hi(n), tile_lo(n), 16 + l, j0, lv);
                        if (lane < 16) { db_at(n % 3, h, lane) = dbm; db_at(n % 3, h, 16 + lane) = dbl; }
                    } else if (q == 0) {
                        // ---- the same with the latent index on lanes 0..L-1 only
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
                                const float m = mu[j] + b_mu, v = lv[j] + b_lv;
                                float gkw = s_wk[jc0 + j];
                                if (PAIRED) {
                                    // d/dKLD of beta w KLD_b + lambda w_o relu(KLD_b - KLD_o + margin): bkg column / OoD column
                                    const int pc = (jc0 + j) & 31;
                                    const bool act = s_kld[pc] - s_kld[pc + 32] + a.margin > 0.f;
                                    gkw = jc0 < 32 ? gkw + (act ? s_wk[pc + 32] : 0.f) : (act ? -gkw : 0.f);
                                }
                                const float s_raw = __expf(v * 0.5f), e_raw = s_raw * s_raw;
                                gm = gz[j] + gkw * m * inv_l;
                                gl = gkw * (-(1.f - (clip_pass(e_raw) ? e_raw : 0.f)) * 0.5f * inv_l)
                                   + gz[j] * s_eps[lane * 64 + jc0 + j] * (clip_pass(s_raw) ? s_raw * 0.5f : 0.f);
                            }
                            mu[j] = gm; lv[j] = gl; dbm += gm; dbl += gl;
                        }
                        if (lane < 16) {
                            store16(tile_hi(n), tile_lo(n), lane, jc0, mu);
                            store16(tile_hi(n), tile_lo(n), 16 + lane, jc0, lv);
                            db_at(n % 3, h, lane) = dbm; db_at(n % 3, h, 16 + lane) = dbl;
                        }
                    }
                    publish(&gready[s - 1]);
                    stamp();
                }
                // ======================================================= stage 0: dW_0^T = G_0 x X, X re-staged from L2
                {
                    float4 ra[2], rb[2];
                    ldx(0, ra); ldx(1, rb);
                    for (int c = 0; c < P.nX; c += 2) {
                        stage_x(ra); ldx(c + 2, ra);
                        if (c + 1 < P.nX) { stage_x(rb); ldx(c + 3, rb); }
                    }
                }
            }
            stamp();
            // =========================================================== per-jet losses (models.py:123-132) and batch sums
            worker_sync();
            if (SCORE) {
                if (tid < nvalid) {
                    const int j = tid;
                    auto q4 = [&](const float* b) { return (b[j] + b[64 + j]) + (b[128 + j] + b[192 + j]); };
                    const long row0 = row0_of();
                    if (!NOMSE && a.s_mse) a.s_mse[row0 + j] = q4(s_rec) / (float)D;
                    if (a.s_mae) a.s_mae[row0 + j] = q4(s_db) / (float)D;
                    if (a.s_kld) a.s_kld[row0 + j] = q4(s_db + 256);
                    if (a.s_jsd) a.s_jsd[row0 + j] = q4(s_db + 512) * 0.5f;
                    if (a.s_lat) {                                                              // latent_loss, utils.py:378-386
                        const float kl = a.heads_t ? -(q4(s_db + 1024) / (float)L) / 2.f : s_kld[j];
                        a.s_lat[row0 + j] = isfinite(kl) ? kl : 0.f;
                    }
                    if (a.s_fin) a.s_fin[row0 + j] = 1;                                         // clip_values leaves no non-finite X_hat
                }
            } else if (tid < 64) {
                const int j = tid;
                float lr_ = 0.f, lk = 0.f, tot = 0.f;
                float lo_ = 0.f;
                if (j < nvalid && (j < 32 || !PAIRED)) {
                    const float rec = (s_rec[j] + s_rec[64 + j] + s_rec[128 + j] + s_rec[192 + j]) / (float)D, wj = s_w[j];
                    lr_ = rec * wj; lk = s_kld[j] * s_wk[j];
                    if (PAIRED) lo_ = fmaxf(s_kld[j] - s_kld[j + 32] + a.margin, 0.f) * s_wk[j + 32];    // lambda w_OoD relu(.)
                    tot = lr_ + lk + lo_;
                    if (a.per_jet) {
                        const long row0 = row0_of();
                        a.per_jet[row0 + j] = lr_; a.per_jet[a.B + row0 + j] = lk;
                        a.per_jet[2 * a.B + row0 + j] = lo_; a.per_jet[3 * a.B + row0 + j] = tot;
                    }
                }
                if (a.sums) {
                    const double v0 = warp_sum((double)lr_), v1 = warp_sum((double)lk), v2 = warp_sum((double)lo_), v3 = warp_sum((double)tot);
                    if (lane == 0) { atomicAdd(a.sums, v0); atomicAdd(a.sums + 1, v1); if (PAIRED) atomicAdd(a.sums + 2, v2); atomicAdd(a.sums + 3, v3); }
                }
            }
        }
    }
    stamp();
    if (a.dbg && tid == 0 && blockIdx.x < 160) {              // per-CTA wall-clock span: a.dbg[512 + 2 b] = entry, [513 + 2 b] = exit (ns)
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        a.dbg[512 + 2 * blockIdx.x] = (long long)dbg_t0; a.dbg[513 + 2 * blockIdx.x] = (long long)t1;
    }
    if (a.dbg && blockIdx.x == 0 && tid == 0) { a.dbg[0] = n_dbg; a.dbg[906] = clock64(); }
    if (a.dbg && blockIdx.x == 0 && warp == kF3Workers && lane == 0) a.dbg[256] = n_dbg2;
    tc_fence_before();
    if (!SCORE && a.grp_ctr && is_train) __threadfence();        // every thread's stores / reductions / atomics are performed before ...
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
    // ... this CTA reports itself finished (EarlyGroups::fin; same unit as the drain warps' counters: 4 per CTA)
    if (!SCORE && a.grp_ctr && is_train && tid == 32) { __threadfence(); atomicAdd(a.grp_ctr + 2, 4u); }
}

}  // namespace oevae
