// liboevae.so: C ABI (include/oevae.h) over the sm_100a kernels.  Host orchestration only; no torch types.
#include "../../include/oevae_test.h"
#include "common.cuh"
#include "ref_kernels.cuh"
#include "prep_hist_kernels.cuh"
#include "roc_kernels.cuh"
#include "scaler_kernels.cuh"
#include "umma_gemm.cuh"
#include "umma_probe.cuh"
#include "engine.h"
#include "prof.h"
#include "adam_kernels.cuh"
#include "fused2_train.cuh"
#include "fused3_train.cuh"
#include "dp_kernels.cuh"
#include "fused2_host.cuh"
#include <cstdlib>
#include <utility>

#include <algorithm>
#include <atomic>
#include <map>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <mutex>
#include <vector>

using namespace oevae;

std::atomic<long long> g_launches{0};
namespace oevae { Profiler g_prof; }
static thread_local std::string g_create_error;

#define LAUNCHED() (g_launches.fetch_add(1, std::memory_order_relaxed))

static inline unsigned cdiv(long a, long b) { return (unsigned)((a + b - 1) / b); }

// ------------------------------------------------------------------------------------------------
int Engine::fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    err = buf;
    return code;
}

int Engine::cuda_check(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    return fail(OEVAE_ECUDA, "%s: %s", what, cudaGetErrorString(e));
}

#define CK(call) do { int _r = cuda_check((call), #call); if (_r) return _r; } while (0)
#define CKL() do { LAUNCHED(); int _r = cuda_check(cudaGetLastError(), "kernel launch"); if (_r) return _r; } while (0)
#define TRY(call) do { int _r = (call); if (_r) return _r; } while (0)

void Knobs::read_env() {
    auto geti = [](const char* k, long long dflt) { const char* e = getenv(k); return e ? atoll(e) : dflt; };
    fused = geti("OEVAE_FUSED", 1) != 0;
    fused_v = (int)geti("OEVAE_FUSED_V", 3);
    oe_layered = geti("OEVAE_OE_LAYERED", 0) != 0;
    score_layered = geti("OEVAE_SCORE_LAYERED", 0) != 0;
    score_static = geti("OEVAE_SCORE_STATIC", 0) != 0;
    score_heads_t = (int)geti("OEVAE_SCORE_HEADS_T", 0);
    fused_dbg = getenv("OEVAE_FUSED_DBG") != nullptr;
    dbg_skip = geti("OEVAE_DBG_SKIP", 0);
    dp_stamps = geti("OEVAE_DP_STAMPS", 0) != 0;
    adam_early = (int)geti("OEVAE_ADAM_EARLY", 27);
}

int Engine::init(const oevae_config& c) {
    cfg = c;
    knobs.read_env();
    if (c.input_dim <= 0 || c.latent_dim <= 0 || c.n_hidden < 0 || c.n_hidden > 8) return fail(OEVAE_EINVAL, "bad layer configuration");
    if (c.oe_type < 0 || c.oe_type > 4) return fail(OEVAE_EINVAL, "bad oe_type %d", c.oe_type);
    if (c.noise_mode < 0 || c.noise_mode > 2) return fail(OEVAE_EINVAL, "bad noise_mode %d", c.noise_mode);
    if (c.precision != OEVAE_PREC_FP32 && c.precision != OEVAE_PREC_BF16X3 && c.precision != OEVAE_PREC_BF16) return fail(OEVAE_EINVAL, "bad precision %d", c.precision);
    D = c.input_dim; L = c.latent_dim; n = c.n_hidden;
    for (int k = 0; k < n; ++k) { if (c.hidden[k] <= 0) return fail(OEVAE_EINVAL, "bad hidden width"); hid.push_back(c.hidden[k]); }
    // Keras trainable_weights order (models.py:22-28,43-48)
    long off = 0; int d = D;
    auto add = [&](int i, int o) { Layer l; l.in = i; l.out = o; l.w = off; off += (long)i * o; l.b = off; off += o; layers.push_back(l); };
    for (int k = 0; k < n; ++k) { add(d, hid[k]); d = hid[k]; }
    add(d, L); add(d, L);
    d = L;
    for (int k = n - 1; k >= 0; --k) { add(d, hid[k]); d = hid[k]; }
    add(d, D);
    P = off;
    maxH = L;
    for (int h : hid) maxH = std::max(maxH, h);
    CK(cudaSetDevice(c.device));
    int cc_major = 0;
    CK(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, c.device));
    if (cc_major != 10) return fail(OEVAE_EUNSUPPORTED, "liboevae is built for sm_100a only (device has compute capability %d.x)", cc_major);
    CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, c.device));
    CK(cudaMalloc(&theta, 3 * P * sizeof(float)));
    m = theta + P; v = m + P;
    CK(cudaMemset(theta, 0, 3 * P * sizeof(float)));
    step = 0; calls = 0;
    Ppad = (P + 3) / 4 * 4;
    TRY(fused_init());
    if (c.max_batch > 0) TRY(reserve(c.max_batch * 2));
    return 0;
}

namespace oevae {
struct FusedState {
    F2Plan plan2{};                                         // feature-major fused kernels (fused2_train.cuh / fused3_train.cuh)
    bool v3 = true;                                         // warp-specialised fused3_kernel (OEVAE_FUSED_V=2 selects fused2, the cross-check)
    unsigned char* wimg = nullptr; int* pos = nullptr; float* parts = nullptr;
    uint32_t img_bytes = 0;
    bool img_dirty = true, last_fused = false;
    int last_grid = 0;
    float* oe_pj = nullptr; float* oe_c = nullptr; long oe_cap = 0;   // fused outlier-exposure step: raw [4][2B] losses, [2][2B] coefficients
    int* tile_ctr = nullptr;                                // dynamic tile counter of this handle's scoring passes (on cfg.device)
    // early parameter groups (adam_kernels.cuh EarlyGroups): counters at tile_ctr[32..33], cumulative target over this handle's fused3
    // training launches (4 drain warps per CTA); grp_ok: the last gradient came from such a launch
    unsigned int grp_total = 0; bool grp_ok = false;
    long long* dbg_buf = nullptr;                           // phase stamps (OEVAE_FUSED_DBG), allocated on demand
};
}

Engine::~Engine() {
    for (int q = 0; q < 8; ++q) if (dp_peer[q] && q != dp_rank) cudaIpcCloseMemHandle(dp_peer[q]);
    if (dp_buf) cudaFree(dp_buf);
    if (theta) cudaFree(theta);
    if (arena) cudaFree(arena);
    if (dp_gsum) cudaFree(dp_gsum);
    if (dp_stamps) cudaFree(dp_stamps);
    if (hl_ring) cudaFree(hl_ring);
    for (int k = 0; k < kHlSlots; ++k) { if (hl_ready[k]) cudaEventDestroy(hl_ready[k]); if (hl_freed[k]) cudaEventDestroy(hl_freed[k]); if (hl_copied[k]) cudaEventDestroy(hl_copied[k]); }
    if (hl_copy) cudaStreamDestroy(hl_copy);
    if (hl_prep) cudaStreamDestroy(hl_prep);
    if (fz) {
        if (fz->wimg) cudaFree(fz->wimg); if (fz->pos) cudaFree(fz->pos); if (fz->parts) cudaFree(fz->parts); if (fz->oe_pj) cudaFree(fz->oe_pj);
        if (fz->oe_c) cudaFree(fz->oe_c); if (fz->tile_ctr) cudaFree(fz->tile_ctr); if (fz->dbg_buf) cudaFree(fz->dbg_buf);
        delete fz;
    }
}

// launch with the programmatic-stream-serialization attribute (PDL): the kernel may start while its predecessor in the
// stream drains; it orders itself with griddepcontrol.wait.  OEVAE_PDL=0 falls back to ordinary launches.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    static const bool pdl = !(getenv("OEVAE_PDL") && atoi(getenv("OEVAE_PDL")) == 0);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// fused kernel set-up: only for bf16x3 and nets that fit one CTA (see build_fused2_plan); otherwise layered path
int Engine::fused_init() {
    if (cfg.precision != OEVAE_PREC_BF16X3 || !knobs.fused) return 0;
    FusedState* f = new FusedState();
    std::vector<int> pos;
    uint32_t img_bytes = 0;
    if (!build_fused2_plan(*this, f->plan2, pos, img_bytes)) { delete f; return 0; }
    fz = f;                                                 // owned by the handle from here on (freed by ~Engine on a failed init too)
    f->img_bytes = img_bytes;
    CK(cudaMalloc(&f->wimg, f->img_bytes));
    CK(cudaMemset(f->wimg, 0, f->img_bytes));
    CK(cudaMalloc(&f->pos, pos.size() * sizeof(int)));
    CK(cudaMemcpy(f->pos, pos.data(), pos.size() * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&f->parts, ((size_t)n_sm * Ppad + 64) * sizeof(float)));
    CK(cudaMalloc(&f->tile_ctr, 256));
    CK(cudaMemset(f->tile_ctr, 0, 256));
    CK(cudaFuncSetAttribute(fused2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    CK(cudaFuncSetAttribute(fused2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    CK(cudaFuncSetAttribute(fused3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    CK(cudaFuncSetAttribute(fused3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    CK(cudaFuncSetAttribute(fused3_kernel<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    CK(cudaFuncSetAttribute(fused3_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    CK(cudaFuncSetAttribute(fused3_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->plan2.smem_bytes));
    f->v3 = knobs.fused_v != 2;
    return 0;
}

// weights were set from the host: rebuild the bf16 images
int Engine::refresh_images(cudaStream_t s) {
    FusedState& f = *fz;
    if (!f.img_dirty) return 0;
    AdamImgArgs ai{};
    ai.theta = theta; ai.n = P; ai.wimg = f.wimg; ai.pos = f.pos;
    { PROF("weight_images", s); adam_img_kernel<<<cdiv(P, kAdamCols), kAdamThreads, 0, s>>>(ai); }
    CKL();
    f.img_dirty = false;
    return 0;
}

long long* Engine::dbg_buffer() {
    if (!knobs.fused_dbg) return nullptr;
    FusedState& f = *fz;
    if (!f.dbg_buf) { cudaMalloc(&f.dbg_buf, 1024 * sizeof(long long)); cudaMemset(f.dbg_buf, 0, 1024 * sizeof(long long)); }
    return f.dbg_buf;
}

int Engine::fused_losses(const float* x, const float* w, const float* eps, long B, float beta, float* per_jet, double* sums,
                         bool want_grad, cudaStream_t s) {
    FusedState& f = *fz;
    TRY(refresh_images(s));
    long long* dbg_buf = dbg_buffer();
    const bool dbg = dbg_buf != nullptr;
    F2Args b{};
    b.x = x; b.w = w; b.eps = eps; b.theta = theta; b.wimg = f.wimg; b.parts = f.parts;
    b.per_jet = per_jet; b.sums = sums;
    b.B = B; b.n_tiles = (int)((B + kF2Jets - 1) / kF2Jets); b.train = want_grad; b.mse = mse_type();
    b.beta = beta; b.seed = cfg.seed; b.row_offset = row_offset;
    b.step = (cfg.noise_mode == OEVAE_NOISE_PER_STEP) ? (unsigned)(4 * calls) : 0u;
    ++calls;
    b.dbg = dbg_buf;
    const int grid2 = std::min(b.n_tiles, n_sm);
    {
        PROF(want_grad ? "fused_train" : "fused_eval", s);
        if (f.v3 && want_grad) b.grp_ctr = reinterpret_cast<unsigned int*>(f.tile_ctr) + 32;
        if (f.v3) CK(launch_pdl(fused3_kernel<false>, dim3(grid2), dim3(kF3Threads), (size_t)f.plan2.smem_bytes, s, f.plan2, b));
        else fused2_kernel<false><<<grid2, kF2Threads, f.plan2.smem_bytes, s>>>(f.plan2, b);
    }
    CKL();
    if (want_grad) { f.last_fused = true; f.last_grid = grid2; f.grp_ok = f.v3; if (f.v3) f.grp_total += 4u * (unsigned)grid2; }
    if (dbg) {
        long long hbuf[1024]; cudaStreamSynchronize(s); cudaMemcpy(hbuf, dbg_buf, sizeof hbuf, cudaMemcpyDeviceToHost);
        fprintf(stderr, "[fused%d dbg] grid=%d stamps(cycles since start):", f.v3 ? 3 : 2, grid2);
        for (int i = 2; i <= (int)hbuf[0]; ++i) fprintf(stderr, " %lld", hbuf[i] - hbuf[1]);
        fprintf(stderr, "\n");
        if (f.v3) {
            long long t0min = 1ll << 62, t1max = 0, dmax = 0, dmin = 1ll << 62; int imax = 0;
            for (int b2 = 0; b2 < grid2 && b2 < 160; ++b2) {
                const long long t0 = hbuf[512 + 2 * b2], t1 = hbuf[513 + 2 * b2];
                t0min = std::min(t0min, t0); t1max = std::max(t1max, t1);
                if (t1 - t0 > dmax) { dmax = t1 - t0; imax = b2; }
                dmin = std::min(dmin, t1 - t0);
            }
            fprintf(stderr, "[fused3 dbg] set-up cycles (tid 0): prefetch %lld, barrier init %lld, zero %lld, tmem alloc %lld, sync %lld; first stamp at +%lld\n",
                    hbuf[901] - hbuf[900], hbuf[902] - hbuf[901], hbuf[903] - hbuf[902], hbuf[904] - hbuf[903], hbuf[905] - hbuf[904], hbuf[1] - hbuf[900]);
            fprintf(stderr, "[fused3 dbg] CTA spans (ns): first entry -> last exit %lld, per-CTA min %lld max %lld (cta %d), cta0 %lld, entry skew %lld\n",
                    t1max - t0min, dmin, dmax, imax, hbuf[513] - hbuf[512], hbuf[512 + 2 * (grid2 - 1)] - t0min);
        }
        if (f.v3 && hbuf[256] > 0) {
            fprintf(stderr, "[fused3 dbg] issuer stamps:");
            for (int i = 0; i < (int)hbuf[256]; ++i) fprintf(stderr, " %lld", hbuf[257 + i] - hbuf[1]);
            fprintf(stderr, "\n");
        }
    }
    return 0;
}

// workspace for `rows` total rows (all passes of one call)
int Engine::reserve(long rows) {
    if (rows <= cap_rows) return 0;
    if (arena) { CK(cudaDeviceSynchronize()); CK(cudaFree(arena)); arena = nullptr; }
    long sumH = 0; for (int h : hid) sumH += h;
    long per_row = 2 * sumH + 6L * L + 2L * D + 2L * maxH + 4;
    Ppad = (P + 3) / 4 * 4;
    size_t floats = (size_t)per_row * rows + (size_t)kMaxSlots * Ppad + 64;
    cudaError_t e = cudaMalloc(&arena, floats * sizeof(float));
    if (e != cudaSuccess) { arena = nullptr; cap_rows = 0; return fail(OEVAE_ENOMEM, "workspace of %zu MB: %s", floats * 4 >> 20, cudaGetErrorString(e)); }
    cap_rows = rows;
    float* p = arena;
    auto take = [&](size_t nflt) { float* q = p; p += (nflt + 3) / 4 * 4; return q; };
    parts = take((size_t)kMaxSlots * Ppad);
    act_e.assign(n + 1, nullptr); act_d.assign(n + 1, nullptr);
    for (int k = 1; k <= n; ++k) act_e[k] = take((size_t)rows * hid[k - 1]);
    mu = take((size_t)rows * L); lv = take((size_t)rows * L);
    z = take((size_t)rows * L); eps_buf = take((size_t)rows * L);
    g_mu = take((size_t)rows * L); g_lv = take((size_t)rows * L);
    for (int k = 1; k <= n; ++k) act_d[k] = take((size_t)rows * hid[n - k]);
    y = take((size_t)rows * D); ysum = take((size_t)rows * D);
    gA = take((size_t)rows * maxH); gB = take((size_t)rows * maxH);
    kld = take(rows); rec = take(rows); g_rec = take(rows); g_kld = take(rows);
    return 0;
}

// ---------------------------------------------------------------------------------------------- GEMM launchers
template <bool SPLIT>
static int umma_smem_bytes(int BN) {                        // two operand stages (MN-major B: whole 64-row blocks) or the epilogue staging tile
    return std::max(2 * (128 * 128 + (BN + 63) / 64 * 8192) * (SPLIT ? 2 : 1), 128 * (BN + 4) * 4) + 1024;
}

template <bool A_TC, bool B_TC, bool SPLIT>
static cudaError_t umma_launch(const GemmArgs& g, int n_slots, int n_sm, cudaStream_t s) {
    int BN = std::min(128, (g.Cn + 15) / 16 * 16);
    int tmem_cols = 32; while (tmem_cols < BN) tmem_cols <<= 1;
    int smem = umma_smem_bytes<SPLIT>(BN);
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(umma_gemm_kernel<A_TC, B_TC, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, umma_smem_bytes<SPLIT>(128));
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const long n_tiles = (long)cdiv(g.R, 128) * cdiv(g.Cn, BN) * n_slots;
    umma_gemm_kernel<A_TC, B_TC, SPLIT><<<(unsigned)std::min<long>(n_tiles, n_sm), kGT, smem, s>>>(g, BN, tmem_cols, n_slots);
    return cudaSuccess;
}

// kind: 0 forward (A t-contiguous, B c-contiguous), 1 dX (both t-contiguous), 2 dW (A r-contiguous, B c-contiguous)
int Engine::launch_gemm(const GemmArgs& g, int kind, int n_slots, cudaStream_t s) {
    static const char* names[3] = {"gemm_fwd", "gemm_dx", "gemm_dw"};
    PROF(names[kind], s);
    if (cfg.precision == OEVAE_PREC_FP32) {
        dim3 grid(cdiv(g.R, 64), cdiv(g.Cn, 64), n_slots);
        if (kind == 0) gemm64_kernel<true, false><<<grid, 256, 0, s>>>(g);
        else if (kind == 1) gemm64_kernel<true, true><<<grid, 256, 0, s>>>(g);
        else gemm64_kernel<false, false><<<grid, 256, 0, s>>>(g);
    } else {
        const bool split = cfg.precision == OEVAE_PREC_BF16X3;
        cudaError_t e;
        if (kind == 0) e = split ? umma_launch<true, false, true>(g, n_slots, n_sm, s) : umma_launch<true, false, false>(g, n_slots, n_sm, s);
        else if (kind == 1) e = split ? umma_launch<true, true, true>(g, n_slots, n_sm, s) : umma_launch<true, true, false>(g, n_slots, n_sm, s);
        else e = split ? umma_launch<false, false, true>(g, n_slots, n_sm, s) : umma_launch<false, false, false>(g, n_slots, n_sm, s);
        CK(e);
    }
    CKL();
    return 0;
}

int Engine::gemm_fwd(const float* A, int li, float* C, long rows, bool relu, cudaStream_t s) {
    const Layer& l = layers[li];
    GemmArgs g{};
    g.A = A; g.B = theta + l.w; g.C = C;
    g.sAr = l.in; g.sAt = 1; g.sBt = l.out; g.sBc = 1;
    g.R = (int)rows; g.Cn = l.out; g.ldc = l.out;
    g.bias = theta + l.b; g.relu = relu;
    g.t_base = 0; g.t_end = l.in;
    return launch_gemm(g, 0, 1, s);
}

// C[rows, in] = G[rows, out] . W^T  (+= C) (* mask>0)
int Engine::gemm_dx(const float* G, int li, float* C, long rows, bool accumulate, const float* mask, cudaStream_t s) {
    const Layer& l = layers[li];
    GemmArgs g{};
    g.A = G; g.B = theta + l.w; g.C = C;
    g.sAr = l.out; g.sAt = 1; g.sBt = 1; g.sBc = l.out;
    g.R = (int)rows; g.Cn = l.in; g.ldc = l.in;
    g.accumulate = accumulate; g.mask = mask; g.ldmask = l.in;
    g.t_base = 0; g.t_end = l.out;
    return launch_gemm(g, 1, 1, s);
}

// parts[slot][w_off + i*out + j] = sum_{rows of slot} [A,1][row][i] * G[row][j]
// A/G point at the first row of the launch; rows are split as n_groups groups of B rows, Sg slots per group.
int Engine::gemm_dw(const float* A, const float* G, int li, long rows, long B, int slot0, int n_slots, cudaStream_t s) {
    const Layer& l = layers[li];
    GemmArgs g{};
    g.A = A; g.B = G; g.C = parts + (size_t)slot0 * Ppad + l.w;
    g.sAr = 1; g.sAt = l.in; g.sBt = l.out; g.sBc = 1;
    g.R = l.in + 1; g.Cn = l.out; g.ldc = l.out;
    g.ones_row = 1;
    g.chunk = chunk_rows; g.slots_per_group = Sg; g.group_rows = (int)B;
    g.t_base = 0; g.t_end = (int)rows; g.slot_stride = Ppad;
    return launch_gemm(g, 2, n_slots, s);
}

// ---------------------------------------------------------------------------------------------- forward
int Engine::forward(const Pass& p, cudaStream_t s) {
    const long B = p.B, Rt = (long)p.R * B, Rd = (long)p.Rd * B;
    // encoder (models.py:30-36)
    const float* a = nullptr;
    for (int k = 0; k < n; ++k) {
        if (k == 0) { for (int g = 0; g < p.R; ++g) TRY(gemm_fwd(p.x[g], 0, act_e[1] + (size_t)g * B * hid[0], B, true, s)); }
        else TRY(gemm_fwd(a, k, act_e[k + 1], Rt, true, s));
        a = act_e[k + 1];
    }
    if (n == 0) {
        for (int g = 0; g < p.R; ++g) {
            TRY(gemm_fwd(p.x[g], 0, mu + (size_t)g * B * L, B, false, s));
            TRY(gemm_fwd(p.x[g], 1, lv + (size_t)g * B * L, B, false, s));
        }
    } else {
        TRY(gemm_fwd(a, n, mu, Rt, false, s));
        TRY(gemm_fwd(a, n + 1, lv, Rt, false, s));
    }
    // Sampling + KLD (models.py:9-15,81-85)
    LatentArgs la{};
    la.mu = mu; la.lv = lv; la.z = z; la.eps_out = eps_buf; la.kld = kld;
    la.eps_ext = p.eps; la.n_eps = p.n_eps; la.n_rows = Rt; la.n_dec = Rd; la.B = B; la.L = L;
    la.seed = cfg.seed; la.row_offset = row_offset;
    la.step_base = (cfg.noise_mode == OEVAE_NOISE_PER_STEP) ? (unsigned)(4 * calls) : 0u;
    la.per_group_step = (cfg.noise_mode == OEVAE_NOISE_PER_STEP);
    { PROF("latent_fwd", s); latent_fwd_kernel<<<cdiv(Rt, 128), 128, 0, s>>>(la); }
    CKL();
    ++calls;
    if (Rd == 0) return 0;
    // decoder (models.py:49-52) + clip + reconstruction loss (models.py:66,72-78)
    const float* d = z;
    for (int k = 0; k < n; ++k) { TRY(gemm_fwd(d, n + 2 + k, act_d[k + 1], Rd, true, s)); d = act_d[k + 1]; }
    TRY(gemm_fwd(d, 2 * n + 2, y, Rd, false, s));
    if (p.want_rec) {
        for (int g = 0; g < p.Rd; ++g) {
            { PROF("rec_fwd", s); rec_fwd_kernel<<<cdiv(B * 32, 256), 256, 0, s>>>(p.x[g], y + (size_t)g * B * D, rec + (size_t)g * B, B, D, mse_type()); }
            CKL();
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------- backward
int Engine::backward(const Pass& p, cudaStream_t s) {
    const long B = p.B, Rt = (long)p.R * B, Rd = (long)p.Rd * B;
    Sg = (int)std::min<long>(kMaxSlots / 3, std::max<long>(1, (B + 255) / 256));
    chunk_rows = (int)((B + Sg - 1) / Sg);
    const int S = Sg * p.R;
    n_slots_used = S;
    float* g = nullptr; float* gn = nullptr;
    // decoder side
    for (int gi = 0; gi < p.Rd; ++gi) {
        { PROF("rec_bwd", s); rec_bwd_kernel<<<cdiv(B * D, 256), 256, 0, s>>>(p.x[gi], y + (size_t)gi * B * D, g_rec + (size_t)gi * B, B, D, mse_type()); }
        CKL();
    }
    g = y;
    for (int k = n; k >= 0; --k) {
        int li = n + 2 + k;
        const float* inp = (k == 0) ? z : act_d[k];
        TRY(gemm_dw(inp, g, li, Rd, B, 0, S, s));
        gn = (g == gA) ? gB : gA;
        TRY(gemm_dx(g, li, gn, Rd, false, k > 0 ? inp : nullptr, s));
        g = gn;
    }
    // g == g_z [Rd, L]
    { PROF("latent_bwd", s); latent_bwd_kernel<<<cdiv(Rt * L, 256), 256, 0, s>>>(mu, lv, eps_buf, g, g_kld, g_mu, g_lv, Rt, Rd, L); }
    CKL();
    if (n == 0) {
        for (int gi = 0; gi < p.R; ++gi) {
            TRY(gemm_dw(p.x[gi], g_mu + (size_t)gi * B * L, 0, B, B, gi * Sg, Sg, s));
            TRY(gemm_dw(p.x[gi], g_lv + (size_t)gi * B * L, 1, B, B, gi * Sg, Sg, s));
        }
        return 0;
    }
    TRY(gemm_dw(act_e[n], g_mu, n, Rt, B, 0, S, s));
    TRY(gemm_dw(act_e[n], g_lv, n + 1, Rt, B, 0, S, s));
    gn = (g == gA) ? gB : gA;
    TRY(gemm_dx(g_mu, n, gn, Rt, false, nullptr, s));
    TRY(gemm_dx(g_lv, n + 1, gn, Rt, true, act_e[n], s));
    g = gn;
    for (int k = n - 1; k >= 0; --k) {
        if (k == 0) {
            for (int gi = 0; gi < p.R; ++gi) TRY(gemm_dw(p.x[gi], g + (size_t)gi * B * hid[0], 0, B, B, gi * Sg, Sg, s));
        } else {
            TRY(gemm_dw(act_e[k], g, k, Rt, B, 0, S, s));
            gn = (g == gA) ? gB : gA;
            TRY(gemm_dx(g, k, gn, Rt, false, act_e[k], s));
            g = gn;
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------- pass planning
int Engine::plan(Pass& p, const float* x_bkg, const float* x_ood, const float* eps, int n_eps, long B) {
    if (B <= 0) return fail(OEVAE_EINVAL, "empty batch");
    if (!x_bkg) return fail(OEVAE_EINVAL, "x_bkg is null");
    p = Pass{};
    p.B = B; p.want_rec = true;
    p.x[0] = x_bkg; p.R = 1; p.Rd = 1;
    if (x_ood) {
        if (cfg.oe_type == OEVAE_OE_KLD) { p.x[1] = x_ood; p.R = 2; p.Rd = 1; }
        else {
            bool fresh = cfg.noise_mode == OEVAE_NOISE_PER_STEP || (cfg.noise_mode == OEVAE_NOISE_EXTERNAL && n_eps == 3);
            if (fresh) { p.x[1] = x_bkg; p.x[2] = x_ood; p.R = 3; p.Rd = 3; }
            else { p.x[1] = x_ood; p.R = 2; p.Rd = 2; }
        }
    }
    if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL) {
        if (!eps) return fail(OEVAE_EINVAL, "noise_mode EXTERNAL needs eps");
        if (n_eps != 1 && n_eps < p.Rd) return fail(OEVAE_EINVAL, "n_eps must be 1 or >= %d", p.Rd);
        p.eps = eps; p.n_eps = n_eps;
    } else if (eps) {
        return fail(OEVAE_EINVAL, "eps given but noise_mode is not EXTERNAL");
    }
    return reserve((long)p.R * B);
}

int Engine::losses(const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood, const float* eps, int n_eps,
                   long B, float beta, float lamb, float margin, float* per_jet, double* sums, bool want_grad, cudaStream_t s) {
    if (!w_bkg) return fail(OEVAE_EINVAL, "w_bkg is null");
    if ((x_ood == nullptr) != (w_ood == nullptr)) return fail(OEVAE_EINVAL, "x_ood and w_ood must be given together");
    if (fz && !x_ood && B > 0 && x_bkg) {
        if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL) {
            if (!eps) return fail(OEVAE_EINVAL, "noise_mode EXTERNAL needs eps");
        } else if (eps) return fail(OEVAE_EINVAL, "eps given but noise_mode is not EXTERNAL");
        return fused_losses(x_bkg, w_bkg, eps, B, beta, per_jet, sums, want_grad, s);
    }
    const bool same_eps = cfg.noise_mode == OEVAE_NOISE_FIXED || (cfg.noise_mode == OEVAE_NOISE_EXTERNAL && n_eps == 1);
    if (fz && fz->v3 && x_ood && B > 0 && x_bkg && same_eps && !knobs.oe_layered) {
        if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL) {
            if (!eps) return fail(OEVAE_EINVAL, "noise_mode EXTERNAL needs eps");
        } else if (eps) return fail(OEVAE_EINVAL, "eps given but noise_mode is not EXTERNAL");
        return fused_oe(x_bkg, w_bkg, x_ood, w_ood, eps, B, beta, lamb, margin, per_jet, sums, want_grad, s);
    }
    if (fz && want_grad) fz->last_fused = false;
    Pass p;
    TRY(plan(p, x_bkg, x_ood, eps, n_eps, B));
    TRY(forward(p, s));
    CombineArgs ca{};
    ca.rec = rec; ca.kld = kld; ca.w = w_bkg; ca.w_ood = w_ood;
    ca.g_rec = want_grad ? g_rec : nullptr; ca.g_kld = want_grad ? g_kld : nullptr;
    ca.per_jet = per_jet; ca.sums = sums;
    ca.B = B; ca.n_groups = p.R; ca.oe_type = cfg.oe_type; ca.have_ood = x_ood != nullptr;
    ca.beta = beta; ca.lamb = lamb; ca.margin = margin;
    { PROF("combine", s); combine_kernel<<<cdiv(B, 256), 256, 0, s>>>(ca); }
    CKL();
    if (want_grad) TRY(backward(p, s));
    return 0;
}

// Outlier-exposure step (models.py:86-132, train_step 147-155) on the fused kernel, for the noise modes in which the bkg and OoD
// passes see the same eps (the reference's fixed-seed Sampling layer; EXTERNAL with one eps block):
//   1. forward-only fused3 pass over the virtual batch [bkg ; OoD] with unit weights -> raw per-jet rec / kld of both halves,
//   2. combine_kernel: per-jet OE loss, per-jet outputs / sums, and d loss / d rec, d loss / d kld of every row,
//   3. fused3 training pass over the same 2B rows with those derivatives as the per-jet coefficients of the two loss terms.
int Engine::fused_oe(const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood, const float* eps, long B,
                     float beta, float lamb, float margin, float* per_jet, double* sums, bool want_grad, cudaStream_t s) {
    FusedState& f = *fz;
    TRY(refresh_images(s));
    if (cfg.oe_type == OEVAE_OE_KLD) {
        // paired tiles: the hinge on KLD(bkg) - KLD(OoD) and its derivative are formed inside the kernel -> one launch
        F2Args b{};
        b.paired = 1; b.x = x_bkg; b.x2 = x_ood; b.w = w_bkg; b.w2 = w_ood; b.eps = eps; b.theta = theta; b.wimg = f.wimg; b.parts = f.parts;
        b.per_jet = per_jet; b.sums = sums;
        b.B = B; b.n_tiles = (int)((B + 31) / 32); b.train = want_grad; b.mse = mse_type();
        b.beta = beta; b.lamb = lamb; b.margin = margin; b.seed = cfg.seed; b.row_offset = row_offset; b.step = 0u;
        ++calls;
        const int grid = std::min(b.n_tiles, n_sm);
        if (want_grad) b.grp_ctr = reinterpret_cast<unsigned int*>(f.tile_ctr) + 32;
        { PROF(want_grad ? "fused_oe_pair_train" : "fused_oe_pair_eval", s); CK(launch_pdl(fused3_kernel<false, 1>, dim3(grid), dim3(kF3Threads), (size_t)f.plan2.smem_bytes, s, f.plan2, b)); }
        CKL();
        if (want_grad) { f.last_fused = true; f.last_grid = grid; f.grp_ok = true; f.grp_total += 4u * (unsigned)grid; }
        return 0;
    }
    const long Bt = 2 * B;
    if (Bt > f.oe_cap) {
        CK(cudaStreamSynchronize(s));
        if (f.oe_pj) cudaFree(f.oe_pj);
        if (f.oe_c) cudaFree(f.oe_c);
        f.oe_pj = f.oe_c = nullptr; f.oe_cap = 0;
        CK(cudaMalloc(&f.oe_pj, 4 * Bt * sizeof(float)));
        CK(cudaMalloc(&f.oe_c, 2 * Bt * sizeof(float)));
        f.oe_cap = Bt;
    }
    F2Args b{};
    b.x = x_bkg; b.x2 = x_ood; b.B1 = B; b.eps = eps; b.theta = theta; b.wimg = f.wimg; b.parts = f.parts;
    b.B = Bt; b.n_tiles = 2 * (int)((B + kF2Jets - 1) / kF2Jets); b.mse = mse_type();   // each half starts at a tile boundary
    b.beta = 1.f; b.seed = cfg.seed; b.row_offset = row_offset; b.step = 0u;
    ++calls;
    const int grid = std::min(b.n_tiles, n_sm);
    b.w = nullptr; b.per_jet = f.oe_pj; b.train = 0;        // raw rec -> oe_pj[0..2B), raw kld -> oe_pj[2B..4B)
    { PROF("fused_oe_fwd", s); CK(launch_pdl(fused3_kernel<false, 2>, dim3(grid), dim3(kF3Threads), (size_t)f.plan2.smem_bytes, s, f.plan2, b)); }
    CKL();
    CombineArgs ca{};
    ca.rec = f.oe_pj; ca.kld = f.oe_pj + Bt; ca.w = w_bkg; ca.w_ood = w_ood;
    ca.g_rec = want_grad ? f.oe_c : nullptr; ca.g_kld = want_grad ? f.oe_c + Bt : nullptr;
    ca.per_jet = per_jet; ca.sums = sums;
    ca.B = B; ca.n_groups = 2; ca.oe_type = cfg.oe_type; ca.have_ood = 1;
    ca.beta = beta; ca.lamb = lamb; ca.margin = margin;
    { PROF("combine", s); combine_kernel<<<cdiv(B, 256), 256, 0, s>>>(ca); }
    CKL();
    if (!want_grad) return 0;
    b.c_rec = f.oe_c; b.c_kld = f.oe_c + Bt; b.per_jet = nullptr; b.sums = nullptr; b.train = 1;
    b.grp_ctr = reinterpret_cast<unsigned int*>(f.tile_ctr) + 32;
    { PROF("fused_oe_train", s); CK(launch_pdl(fused3_kernel<false, 2>, dim3(grid), dim3(kF3Threads), (size_t)f.plan2.smem_bytes, s, f.plan2, b)); }
    CKL();
    f.last_fused = true; f.last_grid = grid; f.grp_ok = true; f.grp_total += 4u * (unsigned)grid;
    return 0;
}

// early parameter groups of the update that follows a fused3 training launch (adam_kernels.cuh); all-null when they do not apply
EarlyGroups Engine::early_groups() const {
    EarlyGroups e{};
    if (!fz || !fz->grp_ok || !knobs.adam_early || n < 1) return e;
    e.ctr = reinterpret_cast<const unsigned int*>(fz->tile_ctr) + 32;
    e.target = fz->grp_total;
    if (knobs.adam_early & 1) { e.lo[0] = layers.back().w; e.hi[0] = P; }                  // dense_output: kernel + bias (the last parameters of the flat vector)
    if (knobs.adam_early & 2) { e.lo[1] = layers[1].w; e.hi[1] = layers.back().w; }        // everything between layer 0 and the output layer
    e.dry = (knobs.adam_early & 4) != 0;
    e.fin = (knobs.adam_early & 16) != 0;
    return e;
}

int Engine::reduce_and_update(float* grad_out, const float* grad_in, bool apply, float lr, cudaStream_t s) {
    AdamArgs a{};
    if (grad_in) { a.grad_in = grad_in; }
    else { a.parts = parts; a.n_parts = n_slots_used; a.part_stride = Ppad; a.grad_out = grad_out; }
    a.theta = theta; a.m = m; a.v = v; a.n = P;
    a.b1 = 0.9f; a.b2 = 0.999f; a.eps = 1e-7f; a.apply = apply;
    if (apply) {
        ++step;
        double t = (double)step;
        a.alpha = (float)((double)lr * std::sqrt(1.0 - std::pow(0.999, t)) / (1.0 - std::pow(0.9, t)));
    }
    if (fz) {                                               // same update + refresh of the bf16 weight images
        AdamImgArgs ai{};
        if (grad_in) ai.grad_in = grad_in;
        else if (fz->last_fused) { ai.parts = fz->parts; ai.n_parts = fz->last_grid; ai.part_stride = Ppad; ai.grad_out = grad_out; ai.early = early_groups(); }
        else { ai.parts = parts; ai.n_parts = n_slots_used; ai.part_stride = Ppad; ai.grad_out = grad_out; }
        ai.theta = theta; ai.m = m; ai.v = v; ai.n = P;
        ai.alpha = a.alpha; ai.b1 = a.b1; ai.b2 = a.b2; ai.eps = a.eps; ai.apply = apply;
        ai.wimg = fz->wimg; ai.pos = fz->pos; ai.trace = SumsTrace{hl_trace_src, hl_trace_dst};
        { PROF("adam", s); CK(launch_pdl(adam_img_kernel, dim3((unsigned)cdiv(P, kAdamCols)), dim3(kAdamThreads), 0, s, ai)); }
        CKL();
        if (apply) fz->img_dirty = false;
        return 0;
    }
    { PROF("adam", s); adam_kernel<<<cdiv(P, 256), 256, 0, s>>>(a); }
    CKL();
    return 0;
}

// ---------------------------------------------------------------------------------------------- data parallel over peer memory
// receive buffers: [2 parities][kDpMaxWorld ranks][slots] of {three sums, sequence tag}, kDpSlots slots per block (dp_kernels.cuh)
static long dp_slots(long P) { return (long)cdiv(P, kAdamCols) * kDpSlots; }
static size_t dp_bytes(long P) { return (size_t)2 * kDpMaxWorld * dp_slots(P) * sizeof(uint4); }

int Engine::dp_export(void* handle_out) {
    if (!handle_out) return fail(OEVAE_EINVAL, "dp_export: null handle");
    if (!dp_buf) {
        dp_nblk = (int)cdiv(P, kAdamCols);
        CK(cudaMalloc(&dp_buf, dp_bytes(P)));
        CK(cudaMemset(dp_buf, 0, dp_bytes(P)));
        CK(cudaDeviceSynchronize());
    }
    CK(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle_out), dp_buf));
    return 0;
}

int Engine::dp_connect(int rank, int world, const void* handles) {
    if (!dp_buf) return fail(OEVAE_EINVAL, "dp_connect before dp_export");
    if (world < 2 || world > kDpMaxWorld || rank < 0 || rank >= world || !handles) return fail(OEVAE_EINVAL, "dp_connect: bad rank/world");
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dp_adam_kernel, kAdamThreads, 0));
    if ((long)per_sm * n_sm < dp_nblk)                      // all blocks must be co-resident (see dp_kernels.cuh)
        return fail(OEVAE_EUNSUPPORTED, "dp: %d blocks exceed the %d resident blocks of this GPU; use the all-reduce path", dp_nblk, per_sm * n_sm);
    const cudaIpcMemHandle_t* hs = reinterpret_cast<const cudaIpcMemHandle_t*>(handles);
    for (int q = 0; q < world; ++q) {
        if (q == rank) { dp_peer[q] = dp_buf; continue; }
        CK(cudaIpcOpenMemHandle(&dp_peer[q], hs[q], cudaIpcMemLazyEnablePeerAccess));
    }
    if (!dp_gsum) CK(cudaMalloc(&dp_gsum, Ppad * sizeof(float)));
    if (knobs.dp_stamps && !dp_stamps) { CK(cudaMalloc(&dp_stamps, (size_t)dp_nblk * kDpStamps * sizeof(long long))); CK(cudaMemset(dp_stamps, 0, (size_t)dp_nblk * kDpStamps * sizeof(long long))); }
    // tags restart at 1: no slot may still carry a tag of an earlier connection (the peers push only after the host-side barrier that
    // follows this call, engine.DeviceVAE.dp_connect)
    CK(cudaMemset(dp_buf, 0, dp_bytes(P)));
    CK(cudaDeviceSynchronize());
    dp_rank = rank; dp_world = world; dp_seq = 0; dp_count = 0;
    return 0;
}

// after a gradient pass: local sum of the partial slices + exchange with the peers + clip_values + Adam, one kernel
// have_grad false: this rank's shard of the batch was empty (it still takes part in the exchange, with a zero gradient)
int Engine::dp_update(float lr, bool have_grad, cudaStream_t s) {
    if (dp_world < 2) return fail(OEVAE_EINVAL, "dp_update: not connected");
    DpArgs a{};
    if (!have_grad) { a.parts = fz ? fz->parts : parts; a.n_parts = 0; }
    else if (fz && fz->last_fused) {
        a.parts = fz->parts; a.n_parts = fz->last_grid;
        a.early = early_groups();
        // output-layer / chain groups in the data-parallel update (bit 3): measured -1.2 us per step at 2 GPUs, -0.7 us at 4, -1.8 us at 8 with the
        // push exchange (they move 56 % of the pushed bytes under the backward pass); with the round-1 flag + pull exchange they cost 2 us
        if (!(knobs.adam_early & 8)) { a.early.lo[0] = a.early.hi[0] = a.early.lo[1] = a.early.hi[1] = -1; }   // only the finished-CTA counter
        if (!a.early.fin) a.early = EarlyGroups{};
    }
    else { a.parts = parts; a.n_parts = n_slots_used; }
    a.gsum_out = dp_gsum; a.stamps = dp_stamps; a.trace = SumsTrace{hl_trace_src, hl_trace_dst};
    a.part_stride = Ppad;
    ++step; ++dp_seq;
    if (dp_seq == 0) ++dp_seq;                              // 0 is the tag of an empty slot
    a.stride = dp_slots(P);
    const size_t par = (size_t)(++dp_count & 1) * kDpMaxWorld * a.stride;    // parity alternates with every exchange of this connection (every rank takes part in each)
    a.my_recv = reinterpret_cast<const uint4*>(dp_buf) + par;
    for (int q = 0; q < dp_world; ++q) a.peer_recv[q] = reinterpret_cast<uint4*>(dp_peer[q]) + par;
    a.rank = dp_rank; a.world = dp_world; a.seq = dp_seq;
    a.theta = theta; a.m = m; a.v = v; a.n = P;
    const double t = (double)step;
    a.alpha = (float)((double)lr * std::sqrt(1.0 - std::pow(0.999, t)) / (1.0 - std::pow(0.9, t)));
    a.b1 = 0.9f; a.b2 = 0.999f; a.eps = 1e-7f;
    if (fz) { a.wimg = fz->wimg; a.pos = fz->pos; }
    { PROF("dp_adam", s); CK(launch_pdl(dp_adam_kernel, dim3((unsigned)dp_nblk), dim3(kAdamThreads), 0, s, a)); }
    CKL();
    if (fz) fz->img_dirty = false;
    return 0;
}

// train_model's batch loop over a host-resident load (models.py:172-190): see oevae.h.  Slot k of the ring (or row b * stride of
// the resident buffers) is filled on the copy stream and consumed on the caller's stream; ready[k] / freed[k] order the two.
int Engine::train_batches_host(oevae_host_load& a, cudaStream_t s) {
    if (!a.x_host || !a.w_host || a.n_rows <= 0 || a.batch_size <= 0 || a.first_batch < 0 || a.n_batches < 0) return fail(OEVAE_EINVAL, "train_batches_host: bad arguments");
    if ((a.xo_host == nullptr) != (a.wo_host == nullptr)) return fail(OEVAE_EINVAL, "train_batches_host: xo_host and wo_host must be given together");
    if (a.x_dtype != 0 && a.x_dtype != 1) return fail(OEVAE_EINVAL, "train_batches_host: x_dtype must be 0 (float32) or 1 (float16)");
    if (a.world < 1 || a.rank < 0 || a.rank >= a.world) return fail(OEVAE_EINVAL, "train_batches_host: bad rank / world");
    if (a.update == 1 && dp_world < 2) return fail(OEVAE_EINVAL, "train_batches_host: update 1 needs oevae_dp_connect");
    if (a.update != 0 && a.update != 1) return fail(OEVAE_EINVAL, "train_batches_host: bad update mode");
    if (a.center && !a.scale) return fail(OEVAE_EINVAL, "train_batches_host: center without scale");
    if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL) return fail(OEVAE_EINVAL, "train_batches_host: noise_mode EXTERNAL has no eps source here");
    const bool ood = a.xo_host != nullptr, f16 = a.x_dtype == 1, resident = a.x_resident != nullptr;
    if (resident && (!a.w_resident || (ood && (!a.xo_resident || !a.wo_resident)))) return fail(OEVAE_EINVAL, "train_batches_host: incomplete resident buffers");
    const long cap = (a.batch_size + a.world - 1) / a.world + 1;                 // largest shard of a global batch
    if (resident && a.resident_stride < cap - 1) return fail(OEVAE_EINVAL, "train_batches_host: resident_stride too small");
    if (!hl_copy) {
        CK(cudaStreamCreateWithFlags(&hl_copy, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&hl_prep, cudaStreamNonBlocking));
        for (int k = 0; k < kHlSlots; ++k) {
            CK(cudaEventCreateWithFlags(&hl_ready[k], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&hl_freed[k], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&hl_copied[k], cudaEventDisableTiming));
        }
    }
    // the per-step read-back of the running sums: stored by the update kernel straight into the caller's page-locked buffer when the
    // device can address it (no copy node between the kernels of consecutive steps), else a 32-byte copy after every step
    double* trace_dev = nullptr;
    const bool trace_direct = a.sums_trace_host && a.sums_acc && fz && cudaHostGetDevicePointer(reinterpret_cast<void**>(&trace_dev), a.sums_trace_host, 0) == cudaSuccess && trace_dev;
    if (!trace_direct) cudaGetLastError();
    // ring slot: [x fp32][xo fp32][w][wo][x f16][xo f16]; with resident buffers only the f16 staging part is used
    const size_t xb = (size_t)cap * D * 4, wb = ((size_t)cap * 4 + 255) / 256 * 256, hb = ((size_t)cap * D * 2 + 255) / 256 * 256;
    const size_t slot_bytes = 2 * xb + 2 * wb + 2 * hb;
    constexpr int NS = kHlSlots, AHEAD = kHlSlots - 1;
    if (hl_ring_bytes < NS * slot_bytes) {
        CK(cudaStreamSynchronize(s)); CK(cudaStreamSynchronize(hl_copy)); CK(cudaStreamSynchronize(hl_prep));
        if (hl_ring) cudaFree(hl_ring);
        hl_ring = nullptr; hl_ring_bytes = 0;
        CK(cudaMalloc(&hl_ring, NS * slot_bytes));
        hl_ring_bytes = NS * slot_bytes;
        for (int k = 0; k < NS; ++k) hl_used[k] = false;
    }
    const size_t esz = f16 ? 2 : 4;
    const long nb_total = (a.n_rows + a.batch_size - 1) / a.batch_size;
    const long b_end = std::min<long>(a.first_batch + a.n_batches, nb_total);
    struct Shard { long lo, n, off; };
    auto shard = [&](long b) {
        const long g0 = b * a.batch_size, g1 = std::min<long>(g0 + a.batch_size, a.n_rows), n = g1 - g0;
        const long s0 = (n * a.rank) / a.world, s1 = (n * (a.rank + 1)) / a.world;
        return Shard{g0 + s0, s1 - s0, s0};
    };
    struct Bufs { float *x, *xo, *w, *wo; unsigned char *hx, *hxo; };
    auto bufs = [&](long b) {
        unsigned char* base = hl_ring + (size_t)(b % NS) * slot_bytes;
        Bufs r;
        r.x = reinterpret_cast<float*>(base); r.xo = reinterpret_cast<float*>(base + xb);
        r.w = reinterpret_cast<float*>(base + 2 * xb); r.wo = reinterpret_cast<float*>(base + 2 * xb + wb);
        r.hx = base + 2 * xb + 2 * wb; r.hxo = r.hx + hb;
        if (resident) {
            r.x = a.x_resident + (size_t)b * a.resident_stride * D; r.w = a.w_resident + (size_t)b * a.resident_stride;
            if (ood) { r.xo = a.xo_resident + (size_t)b * a.resident_stride * D; r.wo = a.wo_resident + (size_t)b * a.resident_stride; }
        }
        return r;
    };
    auto issue_copy = [&](long b) -> int {
        const int k = (int)(b % NS);
        const Shard sh = shard(b);
        const Bufs d = bufs(b);
        if (hl_used[k]) CK(cudaStreamWaitEvent(hl_copy, hl_freed[k], 0));          // the step that read slot k's previous content is done
        if (sh.n > 0) {
            const size_t xbytes = (size_t)sh.n * D * esz;
            CK(cudaMemcpyAsync(f16 ? (void*)d.hx : (void*)d.x, (const unsigned char*)a.x_host + (size_t)sh.lo * D * esz, xbytes, cudaMemcpyHostToDevice, hl_copy));
            CK(cudaMemcpyAsync(d.w, a.w_host + sh.lo, (size_t)sh.n * 4, cudaMemcpyHostToDevice, hl_copy));
            a.h2d_bytes += (int64_t)(xbytes + (size_t)sh.n * 4);
            if (ood) {
                CK(cudaMemcpyAsync(f16 ? (void*)d.hxo : (void*)d.xo, (const unsigned char*)a.xo_host + (size_t)sh.lo * D * esz, xbytes, cudaMemcpyHostToDevice, hl_copy));
                CK(cudaMemcpyAsync(d.wo, a.wo_host + sh.lo, (size_t)sh.n * 4, cudaMemcpyHostToDevice, hl_copy));
                a.h2d_bytes += (int64_t)(xbytes + (size_t)sh.n * 4);
            }
        }
        // float16 -> float32 + scaler (or the float32 scaler pass) on the preparation stream: off the steps' critical path, and the copy
        // engine goes straight on to the next batch
        if (sh.n > 0 && (f16 || a.scale)) {
            CK(cudaEventRecord(hl_copied[k], hl_copy));
            CK(cudaStreamWaitEvent(hl_prep, hl_copied[k], 0));
            const long ne = sh.n * D;
            if (f16) {
                { PROF("ingest_f16", hl_prep); ingest_f16_kernel<<<cdiv((ne + 7) / 8, 256), 256, 0, hl_prep>>>(reinterpret_cast<const __half*>(d.hx), d.x, ne, D, a.center, a.scale); } CKL();
                if (ood) { { PROF("ingest_f16", hl_prep); ingest_f16_kernel<<<cdiv((ne + 7) / 8, 256), 256, 0, hl_prep>>>(reinterpret_cast<const __half*>(d.hxo), d.xo, ne, D, a.center, a.scale); } CKL(); }
            } else {
                { PROF("affine", hl_prep); affine_kernel<<<cdiv(ne, 256), 256, 0, hl_prep>>>(d.x, d.x, sh.n, D, a.center, a.scale); } CKL();
                if (ood) { { PROF("affine", hl_prep); affine_kernel<<<cdiv(ne, 256), 256, 0, hl_prep>>>(d.xo, d.xo, sh.n, D, a.center, a.scale); } CKL(); }
            }
            CK(cudaEventRecord(hl_ready[k], hl_prep));
        } else {
            CK(cudaEventRecord(hl_ready[k], hl_copy));
        }
        return 0;
    };
    a.h2d_bytes = 0;
    const unsigned long long save_off = row_offset;
    // copies run AHEAD batches ahead of the steps, also across calls: a caller that walks a load in groups (train_model reads the
    // running means every 10 batches, models.py:191) finds the first batches of the next group already in flight
    if (hl_src != a.x_host || hl_copied_upto < a.first_batch || hl_copied_upto > a.first_batch + AHEAD) { hl_src = a.x_host; hl_copied_upto = a.first_batch; }
    auto copy_upto = [&](long upto) -> int {                                       // issue the copies of batches [hl_copied_upto, upto)
        for (; hl_copied_upto < std::min<long>(upto, nb_total); ++hl_copied_upto) TRY(issue_copy(hl_copied_upto));
        return 0;
    };
    TRY(copy_upto(a.first_batch + AHEAD));
    if (a.first_batch < b_end) CK(cudaStreamWaitEvent(s, hl_ready[a.first_batch % NS], 0));
    for (long b = a.first_batch; b < b_end; ++b) {
        const int k = (int)(b % NS);
        const Shard sh = shard(b);
        const Bufs d = bufs(b);
        row_offset = save_off + (unsigned long long)(a.noise_row0 + sh.off);        // noise of a jet does not depend on the sharding
        int rc = 0;
        if (sh.n > 0) rc = losses(d.x, d.w, ood ? d.xo : nullptr, ood ? d.wo : nullptr, nullptr, 0, sh.n, a.beta, a.lamb, a.margin, nullptr, a.sums_acc, true, s);
        // the NEXT batch's data is awaited before this step's update kernel, so that update kernel and next training kernel stay
        // adjacent in the stream (programmatic dependent launch: the training kernel's set-up overlaps the update)
        if (!rc && b + 1 < b_end) CK(cudaStreamWaitEvent(s, hl_ready[(b + 1) % NS], 0));
        if (trace_direct) { hl_trace_src = a.sums_acc; hl_trace_dst = trace_dev + 4 * (b - a.first_batch); }
        if (!rc) rc = a.update == 1 ? dp_update(a.lr, sh.n > 0, s) : (sh.n > 0 ? reduce_and_update(nullptr, nullptr, true, a.lr, s) : 0);
        hl_trace_src = nullptr; hl_trace_dst = nullptr;
        row_offset = save_off;
        if (rc) return rc;
        CK(cudaEventRecord(hl_freed[k], s)); hl_used[k] = true;
        if (a.sums_trace_host && a.sums_acc && (!trace_direct || (a.update == 0 && sh.n == 0)))
            CK(cudaMemcpyAsync(a.sums_trace_host + 4 * (b - a.first_batch), a.sums_acc, 4 * sizeof(double), cudaMemcpyDeviceToHost, s));
        TRY(copy_upto(b + 1 + AHEAD));
    }
    if (b_end >= nb_total) hl_src = nullptr;                                       // load finished: the next call starts afresh
    return 0;
}

// fresh optimiser: tf.keras.optimizers.Adam(lr) is constructed anew by every train_model call (models.py:138)
int Engine::reset_opt() {
    CK(cudaMemset(m, 0, 2 * P * sizeof(float)));
    step = 0;
    return 0;
}

int Engine::score(const float* x, long N, int n_iter, const float* eps, float* o_mse, float* o_mae, float* o_kld, float* o_jsd,
                  float* o_lat, unsigned char* o_fin, float* o_xhat, cudaStream_t s) {
    if (N < 0 || n_iter < 1) return fail(OEVAE_EINVAL, "bad N / n_iter");
    if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL && !eps) return fail(OEVAE_EINVAL, "noise_mode EXTERNAL needs eps");
    if (N == 0) return 0;
    // single sampling pass: the whole pass is one launch of the fused forward kernel in its
    // SCORE instantiation (X read once, nothing but the [N] score vectors written)
    if (fz && n_iter == 1 && !knobs.score_layered) {
        FusedState& f = *fz;
        TRY(refresh_images(s));
        F2Args b{};
        b.x = x; b.w = nullptr; b.eps = eps; b.theta = theta; b.wimg = f.wimg; b.parts = nullptr;
        b.B = N; b.n_tiles = (int)((N + kF2Jets - 1) / kF2Jets); b.train = 0; b.mse = 0; b.beta = 0.f;
        b.seed = cfg.seed; b.row_offset = row_offset;
        b.step = (cfg.noise_mode == OEVAE_NOISE_PER_STEP) ? (unsigned)(4 * calls) : 0u;
        ++calls;
        b.s_mse = o_mse; b.s_mae = o_mae; b.s_kld = o_kld; b.s_jsd = o_jsd; b.s_lat = o_lat; b.s_fin = o_fin; b.s_xhat = o_xhat;
        b.heads_t = knobs.score_heads_t;                     // 1: heads with the jets on the lanes (A/B timing: +5% with two metrics, -3% with four)
        const int grid2 = std::min(b.n_tiles, n_sm);
        // dynamic tile counter of the pass: per handle, on the handle's device, zeroed in stream order (passes of one handle are
        // serialised by its stream; a handle is not thread-safe, oevae.h)
        CK(cudaMemsetAsync(f.tile_ctr, 0, sizeof(int), s));
        b.tile_ctr = knobs.score_static ? nullptr : f.tile_ctr;   // OEVAE_SCORE_STATIC=1: round-robin tiles (A/B timing, bit-identical scores)
        long long* sdbg = dbg_buffer();
        const bool dbg = sdbg != nullptr;
        b.dbg = sdbg;
        if (dbg) { long long skip = knobs.dbg_skip; cudaMemcpy(sdbg + 1000, &skip, sizeof skip, cudaMemcpyHostToDevice); }
        {
            PROF("fused_score", s);
            // workers + issuer + producer + X stagers; the instantiation without the MSE vector keeps 16 more registers for the rest
            if (f.v3 && o_mse) fused3_kernel<true><<<grid2, kF3ScoreThreads, f.plan2.smem_bytes, s>>>(f.plan2, b);
            else if (f.v3) fused3_kernel<true, 3><<<grid2, kF3ScoreThreads, f.plan2.smem_bytes, s>>>(f.plan2, b);
            else fused2_kernel<true><<<grid2, kF2Threads, f.plan2.smem_bytes, s>>>(f.plan2, b);
        }
        CKL();
        if (dbg) {
            long long hbuf[1024]; cudaStreamSynchronize(s); cudaMemcpy(hbuf, sdbg, sizeof hbuf, cudaMemcpyDeviceToHost);
            fprintf(stderr, "[score dbg] grid=%d worker stamps:", grid2);
            for (int i = 2; i <= (int)hbuf[0]; ++i) fprintf(stderr, " %lld", hbuf[i] - hbuf[1]);
            {
                std::vector<std::pair<long long, int>> sp;
                for (int b2 = 0; b2 < grid2 && b2 < 160; ++b2) sp.push_back({hbuf[513 + 2 * b2] - hbuf[512 + 2 * b2], b2});
                std::sort(sp.begin(), sp.end());
                fprintf(stderr, "\n[score dbg] CTA spans (ns): min %lld (cta %d) p25 %lld median %lld p75 %lld max %lld (cta %d); slowest:", sp[0].first, sp[0].second,
                        sp[sp.size() / 4].first, sp[sp.size() / 2].first, sp[3 * sp.size() / 4].first, sp.back().first, sp.back().second);
                for (size_t i = sp.size() >= 8 ? sp.size() - 8 : 0; i < sp.size(); ++i) fprintf(stderr, " %d:%lld", sp[i].second, sp[i].first);
                fprintf(stderr, "; cta0 %lld ns = %lld SM cycles (%.0f MHz)", hbuf[513] - hbuf[512], hbuf[906] - hbuf[900],
                        1e3 * (double)(hbuf[906] - hbuf[900]) / (double)(hbuf[513] - hbuf[512]));
            }
            fprintf(stderr, "\n[score dbg] issuer stamps:");
            for (int i = 0; i < (int)hbuf[256]; ++i) fprintf(stderr, " %lld", hbuf[257 + i] - hbuf[1]);
            fprintf(stderr, "\n");
            fprintf(stderr, "[score dbg] exact-path thread-chunks: %lld of %lld (sum y^2 >= 1e12: %lld, sum x^2 >= 1e12: %lld, min|y| <= 1e-30: %lld, min|x+y| <= 4e-30: %lld)\n",
                    hbuf[950], (long long)((N + 63) / 64) * 512 * f.plan2.nC, hbuf[951], hbuf[952], hbuf[953], hbuf[954]);
            cudaMemset(sdbg + 950, 0, 5 * sizeof(long long));
            fprintf(stderr, "[score dbg] stager stamps (tile start, then one per X chunk):");
            for (int i = 0; i < (int)hbuf[832] && i < 66; ++i) fprintf(stderr, " %lld", hbuf[833 + i] - hbuf[1]);
            fprintf(stderr, "\n");
        }
        return 0;
    }
    long chunk = std::min<long>(N, std::max<long>(cap_rows, 65536));
    TRY(reserve(chunk));
    for (long r0 = 0; r0 < N; r0 += chunk) {
        long nr = std::min(chunk, N - r0);
        for (int it = 0; it < n_iter; ++it) {
            Pass p{};
            p.B = nr; p.R = 1; p.Rd = 1; p.x[0] = x + (size_t)r0 * D; p.want_rec = false;
            if (eps) { p.eps = eps + ((size_t)it * N + r0) * L; p.n_eps = 1; }
            unsigned long long save = row_offset; row_offset += (unsigned long long)r0;
            int rc = forward(p, s);
            row_offset = save;
            if (rc) return rc;
            if (n_iter > 1) { { PROF("accum_clip", s); accum_clip_kernel<<<cdiv(nr * D, 256), 256, 0, s>>>(y, ysum, nr * D, it == 0); } CKL(); }
        }
        bool need = o_mse || o_mae || o_kld || o_jsd || o_fin || o_xhat;
        if (need) {
            MetricOut mo{};
            mo.mse = o_mse ? o_mse + r0 : nullptr; mo.mae = o_mae ? o_mae + r0 : nullptr;
            mo.kld = o_kld ? o_kld + r0 : nullptr; mo.jsd = o_jsd ? o_jsd + r0 : nullptr;
            mo.finite = o_fin ? o_fin + r0 : nullptr; mo.xhat = o_xhat ? o_xhat + (size_t)r0 * D : nullptr;
            { PROF("score_metrics", s); score_metrics_kernel<<<cdiv(nr * 32, 256), 256, 0, s>>>(x + (size_t)r0 * D, n_iter > 1 ? ysum : y, 1.f / (float)n_iter, n_iter == 1, nr, D, mo); }
            CKL();
        }
        if (o_lat) { { PROF("latent_out", s); latent_out_kernel<<<cdiv(nr, 256), 256, 0, s>>>(kld, o_lat + r0, nr); } CKL(); }
    }
    return 0;
}

int Engine::encode(const float* x, long N, const float* eps, float* o_mu, float* o_lv, float* o_z, cudaStream_t s) {
    if (N <= 0) return fail(OEVAE_EINVAL, "bad N");
    if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL && !eps && o_z) return fail(OEVAE_EINVAL, "noise_mode EXTERNAL needs eps");
    long chunk = std::min<long>(N, std::max<long>(cap_rows, 65536));
    TRY(reserve(chunk));
    for (long r0 = 0; r0 < N; r0 += chunk) {
        long nr = std::min(chunk, N - r0);
        Pass p{};
        p.B = nr; p.R = 1; p.Rd = 1; p.x[0] = x + (size_t)r0 * D; p.want_rec = false; p.encoder_only = true;
        if (eps) { p.eps = eps + (size_t)r0 * L; p.n_eps = 1; }
        unsigned long long save = row_offset; row_offset += (unsigned long long)r0;
        int rc = forward_encoder(p, s);
        row_offset = save;
        if (rc) return rc;
        if (o_mu) CK(cudaMemcpyAsync(o_mu + (size_t)r0 * L, mu, (size_t)nr * L * 4, cudaMemcpyDeviceToDevice, s));
        if (o_lv) CK(cudaMemcpyAsync(o_lv + (size_t)r0 * L, lv, (size_t)nr * L * 4, cudaMemcpyDeviceToDevice, s));
        if (o_z) CK(cudaMemcpyAsync(o_z + (size_t)r0 * L, z, (size_t)nr * L * 4, cudaMemcpyDeviceToDevice, s));
    }
    return 0;
}

// encoder + sampling only (vae.encoder(X), models.py:30-36)
int Engine::forward_encoder(const Pass& p, cudaStream_t s) {
    const long B = p.B;
    const float* a = p.x[0];
    for (int k = 0; k < n; ++k) { TRY(gemm_fwd(a, k, act_e[k + 1], B, true, s)); a = act_e[k + 1]; }
    TRY(gemm_fwd(a, n, mu, B, false, s));
    TRY(gemm_fwd(a, n + 1, lv, B, false, s));
    LatentArgs la{};
    la.mu = mu; la.lv = lv; la.z = z; la.eps_out = eps_buf; la.kld = kld;
    la.eps_ext = p.eps; la.n_eps = 1; la.n_rows = B; la.n_dec = B; la.B = B; la.L = L;
    la.seed = cfg.seed; la.row_offset = row_offset;
    la.step_base = (cfg.noise_mode == OEVAE_NOISE_PER_STEP) ? (unsigned)(4 * calls) : 0u;
    la.per_group_step = 0;
    if (cfg.noise_mode == OEVAE_NOISE_EXTERNAL && !p.eps) { la.n_dec = 0; }
    { PROF("latent_fwd", s); latent_fwd_kernel<<<cdiv(B, 128), 128, 0, s>>>(la); }
    CKL();
    ++calls;
    return 0;
}

// ---- weighted ROC + cut thresholds (roc_kernels.cuh)
namespace {
struct RocWs {
    unsigned char* base = nullptr; size_t off = 0, cap = 0;
    template <typename T> T* take(size_t n) { off = (off + 255) / 256 * 256; T* p = reinterpret_cast<T*>(base + off); off += n * sizeof(T); return p; }
};
// exclusive scan of in[n] -> out[n] (may alias), total -> *total_dev; tiles: scratch of cdiv(n, kTile) + 1 entries
int roc_scan(const uint32_t* in, uint32_t* out, long n, uint32_t* tiles, uint32_t* total_dev, cudaStream_t s) {
    using namespace roc;
    const unsigned nt = cdiv(n, kTile);
    scan_tiles_kernel<<<nt, kThreads, 0, s>>>(in, n, tiles); LAUNCHED();
    scan_block_kernel<<<1, 1024, 0, s>>>(tiles, nt, total_dev); LAUNCHED();
    scan_apply_kernel<<<nt, kThreads, 0, s>>>(in, n, tiles, out); LAUNCHED();
    return 0;
}
// stable LSD radix sort of (key, value) pairs by key, ascending; the sorted pairs end up in (k0, v0) (four ping-pong passes)
void roc_sort_pairs(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, long n, uint32_t* hist, uint32_t* tiles, cudaStream_t s) {
    using namespace roc;
    const unsigned nt = cdiv(n, kTile);
    uint32_t *kin = k0, *kout = k1, *vin = v0, *vout = v1;
    for (int shift = 0; shift < 32; shift += 8) {
        rs_hist_kernel<<<nt, kThreads, 0, s>>>(kin, n, shift, hist, (int)nt); LAUNCHED();
        roc_scan(hist, hist, 256L * nt, tiles, nullptr, s);
        rs_scatter_kernel<<<nt, kThreads, 0, s>>>(kin, vin, kout, vout, n, shift, hist, (int)nt); LAUNCHED();
        std::swap(kin, kout); std::swap(vin, vout);
    }
}
}  // namespace

// Allocation-free variant of oevae_argsort_f32 (no counterpart in the reference; added for
// callers that may not call cudaMalloc: graph capture, tight serving loops).  Its radix-sort workspace is the file-scope array
// g_sort_scratch: every GPU that loads the library holds kSortScratchBytes of it until exit.  In HOST code the name g_sort_scratch is
// only a stand-in for the symbol, not a device address, so the address each device maps it at is looked up once per device with
// cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.  Nothing in the workspace has to start at
// zero: keys / payloads [0, N), the 256 x tiles digit histogram and the tile sums are all fully rewritten by the kernels of each call
// before they are read.  All calls share the one array per device: two calls that overlap on the GPU (different streams) race.
namespace {
constexpr long kSortStaticMax = OEVAE_ARGSORT_STATIC_MAX;
constexpr size_t kSortScratchBytes = (size_t)kSortStaticMax * 16                         /* keys + payloads, ping and pong */
                                   + (size_t)256 * (kSortStaticMax / roc::kTile) * 4   /* digit histogram: 256 x tiles */
                                   + 4096;                                             /* tile sums of the histogram scan + alignment slack */
static_assert(kSortStaticMax % roc::kTile == 0, "the static maximum is a whole number of tiles");
}
static __device__ __align__(256) unsigned char g_sort_scratch[kSortScratchBytes];
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_sort_scratch_mu;
unsigned char* g_sort_scratch_addr[kMaxDevices] = {};
// device address of g_sort_scratch on `device` (which the caller has made current); looked up once per device
int sort_scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) { g_create_error = "argsort_static: device index out of range"; return OEVAE_EINVAL; }
    std::lock_guard<std::mutex> lk(g_sort_scratch_mu);
    if (!g_sort_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_sort_scratch);
        if (e != cudaSuccess || !p) { g_create_error = std::string("argsort_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e); return OEVAE_ECUDA; }
        g_sort_scratch_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_sort_scratch_addr[device];
    return 0;
}
}  // namespace

// ================================================================================================ C ABI
#define H(h) if (!(h)) return OEVAE_EINVAL; Engine* e = reinterpret_cast<Engine*>(h); cudaSetDevice(e->cfg.device)

extern "C" {

int oevae_version(void) { return OEVAE_VERSION; }

int oevae_create(const oevae_config* cfg, oevae_t* out) {
    if (!cfg || !out) { g_create_error = "null argument"; return OEVAE_EINVAL; }
    Engine* e = new (std::nothrow) Engine();
    if (!e) { g_create_error = "out of host memory"; return OEVAE_ENOMEM; }
    int rc = e->init(*cfg);
    if (rc) { g_create_error = e->err; delete e; *out = nullptr; return rc; }
    *out = reinterpret_cast<oevae_t>(e);
    return 0;
}

int oevae_destroy(oevae_t h) {
    if (!h) return OEVAE_EINVAL;
    Engine* e = reinterpret_cast<Engine*>(h);
    cudaSetDevice(e->cfg.device);
    cudaDeviceSynchronize();
    delete e;
    return 0;
}

const char* oevae_last_error(oevae_t h) {
    if (!h) return g_create_error.c_str();
    return reinterpret_cast<Engine*>(h)->err.c_str();
}

int64_t oevae_n_params(oevae_t h) { return h ? reinterpret_cast<Engine*>(h)->P : -1; }

int oevae_param_layout(oevae_t h, int64_t* offsets, int32_t* dims_in, int32_t* dims_out, int32_t cap) {
    if (!h) return OEVAE_EINVAL;
    Engine* e = reinterpret_cast<Engine*>(h);
    int nl = (int)e->layers.size();
    for (int i = 0; i < nl && i < cap; ++i) {
        if (offsets) { offsets[2 * i] = e->layers[i].w; offsets[2 * i + 1] = e->layers[i].b; }
        if (dims_in) dims_in[i] = e->layers[i].in;
        if (dims_out) dims_out[i] = e->layers[i].out;
    }
    return nl;
}

int oevae_set_params_host(oevae_t h, const float* flat, int64_t nn) {
    H(h);
    if (!flat || nn != e->P) return e->fail(OEVAE_EINVAL, "set_params: expected %ld floats, got %ld", e->P, (long)nn);
    if (e->fz) e->fz->img_dirty = true;
    return e->cuda_check(cudaMemcpy(e->theta, flat, nn * 4, cudaMemcpyHostToDevice), "set_params");
}
int oevae_get_params_host(oevae_t h, float* flat, int64_t nn) {
    H(h);
    if (!flat || nn != e->P) return e->fail(OEVAE_EINVAL, "get_params: expected %ld floats, got %ld", e->P, (long)nn);
    return e->cuda_check(cudaMemcpy(flat, e->theta, nn * 4, cudaMemcpyDeviceToHost), "get_params");
}
int oevae_set_opt_state_host(oevae_t h, const float* mh, const float* vh, int64_t nn, int64_t step) {
    H(h);
    if (!mh || !vh || nn != e->P || step < 0) return e->fail(OEVAE_EINVAL, "set_opt_state: bad arguments");
    int rc = e->cuda_check(cudaMemcpy(e->m, mh, nn * 4, cudaMemcpyHostToDevice), "set_opt_state m");
    if (rc) return rc;
    rc = e->cuda_check(cudaMemcpy(e->v, vh, nn * 4, cudaMemcpyHostToDevice), "set_opt_state v");
    if (rc) return rc;
    e->step = step;
    return 0;
}
int oevae_get_opt_state_host(oevae_t h, float* mh, float* vh, int64_t nn, int64_t* step) {
    H(h);
    if (nn != e->P) return e->fail(OEVAE_EINVAL, "get_opt_state: bad size");
    if (mh) { int rc = e->cuda_check(cudaMemcpy(mh, e->m, nn * 4, cudaMemcpyDeviceToHost), "get_opt_state m"); if (rc) return rc; }
    if (vh) { int rc = e->cuda_check(cudaMemcpy(vh, e->v, nn * 4, cudaMemcpyDeviceToHost), "get_opt_state v"); if (rc) return rc; }
    if (step) *step = e->step;
    return 0;
}
float* oevae_params_device(oevae_t h) { return h ? reinterpret_cast<Engine*>(h)->theta : nullptr; }

int oevae_set_noise_offset(oevae_t h, uint64_t row_offset) { H(h); e->row_offset = row_offset; return 0; }
int oevae_set_oe_type(oevae_t h, int32_t t) { H(h); if (t < 0 || t > 4) return e->fail(OEVAE_EINVAL, "bad oe_type %d", t); e->cfg.oe_type = t; return 0; }

int oevae_train_step(oevae_t h, const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood,
                     const float* eps, int32_t n_eps, int64_t B, float lr, float beta, float lamb, float margin,
                     float* per_jet_out, double* sums_acc, void* stream) {
    H(h);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = e->losses(x_bkg, w_bkg, x_ood, w_ood, eps, n_eps, B, beta, lamb, margin, per_jet_out, sums_acc, true, s);
    if (rc) return rc;
    return e->reduce_and_update(nullptr, nullptr, true, lr, s);
}

int oevae_dp_last_grad_host(oevae_t h, float* grad_host, int64_t nn) {
    H(h);
    if (!grad_host || nn != e->P) return e->fail(OEVAE_EINVAL, "dp_last_grad: expected %ld floats", e->P);
    if (!e->dp_gsum) return e->fail(OEVAE_EINVAL, "dp_last_grad: not connected");
    return e->cuda_check(cudaMemcpy(grad_host, e->dp_gsum, nn * 4, cudaMemcpyDeviceToHost), "dp_last_grad");
}

int oevae_dp_stamps_host(oevae_t h, int64_t* out_host, int64_t cap) {
    H(h);
    if (!e->dp_stamps) return e->fail(OEVAE_EINVAL, "dp_stamps: set OEVAE_DP_STAMPS=1 before oevae_create and connect first");
    const int64_t nn = std::min<int64_t>(cap, (int64_t)e->dp_nblk * kDpStamps);
    if (!out_host || nn <= 0) return e->fail(OEVAE_EINVAL, "dp_stamps: bad arguments");
    int rc = e->cuda_check(cudaMemcpy(out_host, e->dp_stamps, nn * sizeof(long long), cudaMemcpyDeviceToHost), "dp_stamps");
    return rc ? rc : (int)(nn / kDpStamps);
}

int oevae_reset_opt_state(oevae_t h) { H(h); return e->reset_opt(); }

int oevae_dp_export(oevae_t h, void* handle_out64) { H(h); return e->dp_export(handle_out64); }
int oevae_dp_connect(oevae_t h, int32_t rank, int32_t world, const void* handles) { H(h); return e->dp_connect(rank, world, handles); }
int oevae_train_step_dp(oevae_t h, const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood,
                        const float* eps, int32_t n_eps, int64_t B, float lr, float beta, float lamb, float margin,
                        float* per_jet_out, double* sums_acc, void* stream) {
    H(h);
    cudaStream_t s = (cudaStream_t)stream;
    if (e->dp_world < 2) return e->fail(OEVAE_EINVAL, "train_step_dp: call oevae_dp_export / oevae_dp_connect first");
    // B == 0: a rank whose shard of the (short, last) global batch is empty still exchanges -- a zero gradient
    if (B > 0) {
        int rc = e->losses(x_bkg, w_bkg, x_ood, w_ood, eps, n_eps, B, beta, lamb, margin, per_jet_out, sums_acc, true, s);
        if (rc) return rc;
    } else if (B < 0) return e->fail(OEVAE_EINVAL, "negative batch");
    return e->dp_update(lr, B > 0, s);
}

int oevae_loss_and_grad(oevae_t h, const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood,
                        const float* eps, int32_t n_eps, int64_t B, float beta, float lamb, float margin,
                        float* per_jet_out, double* sums_acc, float* grad_out, void* stream) {
    H(h);
    if (!grad_out) return e->fail(OEVAE_EINVAL, "grad_out is null");
    cudaStream_t s = (cudaStream_t)stream;
    if (B == 0) return e->cuda_check(cudaMemsetAsync(grad_out, 0, e->P * sizeof(float), s), "zero gradient of an empty shard");
    int rc = e->losses(x_bkg, w_bkg, x_ood, w_ood, eps, n_eps, B, beta, lamb, margin, per_jet_out, sums_acc, true, s);
    if (rc) return rc;
    return e->reduce_and_update(grad_out, nullptr, false, 0.f, s);
}

int oevae_train_batches_host(oevae_t h, oevae_host_load* load, void* stream) {
    H(h);
    if (!load) return e->fail(OEVAE_EINVAL, "train_batches_host: null load");
    return e->train_batches_host(*load, (cudaStream_t)stream);
}

int oevae_apply_update(oevae_t h, const float* grad, float lr, void* stream) {
    H(h);
    if (!grad) return e->fail(OEVAE_EINVAL, "grad is null");
    return e->reduce_and_update(nullptr, grad, true, lr, (cudaStream_t)stream);
}

int oevae_eval_losses(oevae_t h, const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood,
                      const float* eps, int32_t n_eps, int64_t B, float beta, float lamb, float margin,
                      float* per_jet_out, double* sums_acc, void* stream) {
    H(h);
    return e->losses(x_bkg, w_bkg, x_ood, w_ood, eps, n_eps, B, beta, lamb, margin, per_jet_out, sums_acc, false, (cudaStream_t)stream);
}

int oevae_encode(oevae_t h, const float* x, int64_t N, const float* eps, float* mu, float* log_var, float* z, void* stream) {
    H(h);
    if (!x) return e->fail(OEVAE_EINVAL, "x is null");
    return e->encode(x, N, eps, mu, log_var, z, (cudaStream_t)stream);
}

int oevae_score(oevae_t h, const float* x, int64_t N, int32_t n_iter, const float* eps, float* mse, float* mae, float* kld,
                float* jsd, float* latent, uint8_t* finite, float* xhat, void* stream) {
    H(h);
    if (!x && N > 0) return e->fail(OEVAE_EINVAL, "x is null");
    return e->score(x, N, n_iter, eps, mse, mae, kld, jsd, latent, finite, xhat, (cudaStream_t)stream);
}

// ---- handle-free kernels
static thread_local std::string g_free_error;
static int free_fail(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    g_create_error = std::string(what) + ": " + cudaGetErrorString(e);
    return OEVAE_ECUDA;
}

int oevae_preprocess(const void* raw, int32_t raw_dtype, int64_t N, int32_t n_raw, int32_t n_const, int32_t n_dims,
                     const double* center, const double* scale, int32_t sort, float* x_out, float* pt_out, float* m_out,
                     int32_t device, void* stream) {
    if (N == 0) return 0;
    if (!raw || !x_out || N < 0 || n_raw <= 0 || n_const <= 0 || (n_dims != 3 && n_dims != 4) || (raw_dtype != 0 && raw_dtype != 1) || (center && !scale)) {
        g_create_error = "preprocess: bad arguments"; return OEVAE_EINVAL;
    }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    int n_pad = 32; while (n_pad < n_raw) n_pad <<= 1;
    size_t per_warp = (size_t)n_pad * 8 + (size_t)n_raw * 16 + 16;          // must match preprocess_kernel
    int warps = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)(200 * 1024) / per_warp));
    if (per_warp > 200 * 1024) { g_create_error = "preprocess: n_raw too large"; return OEVAE_EUNSUPPORTED; }
    size_t smem = per_warp * warps;
    cudaStream_t s = (cudaStream_t)stream;
    unsigned grid = cdiv(N, warps);
    if (raw_dtype == 0) {
        if (smem > 48 * 1024) { rc = free_fail(cudaFuncSetAttribute(preprocess_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "smem attr"); if (rc) return rc; }
        preprocess_kernel<float><<<grid, warps * 32, smem, s>>>((const float*)raw, N, n_raw, n_pad, n_const, n_dims, center, scale, sort, x_out, pt_out, m_out);
    } else {
        if (smem > 48 * 1024) { rc = free_fail(cudaFuncSetAttribute(preprocess_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "smem attr"); if (rc) return rc; }
        preprocess_kernel<__half><<<grid, warps * 32, smem, s>>>((const __half*)raw, N, n_raw, n_pad, n_const, n_dims, center, scale, sort, x_out, pt_out, m_out);
    }
    LAUNCHED();
    return free_fail(cudaGetLastError(), "preprocess launch");
}

int oevae_affine(const float* x, float* y, int64_t N, int32_t D, const double* center, const double* scale, int32_t device, void* stream) {
    if (N == 0) return 0;
    if (!x || !y || N < 0 || D <= 0 || !scale) { g_create_error = "affine: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    affine_kernel<<<cdiv(N * D, 256), 256, 0, (cudaStream_t)stream>>>(x, y, N, D, center, scale);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "affine launch");
}

int oevae_ingest_f16(const void* raw_f16, float* x, int64_t N, int32_t D, const double* center, const double* scale, int32_t device, void* stream) {
    if (N == 0) return 0;
    if (!raw_f16 || !x || N < 0 || D <= 0 || (center && !scale)) { g_create_error = "ingest_f16: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    const long n = (long)N * D;
    { PROF("ingest_f16", (cudaStream_t)stream); ingest_f16_kernel<<<cdiv((n + 7) / 8, 256), 256, 0, (cudaStream_t)stream>>>((const __half*)raw_f16, x, n, D, center, scale); }
    LAUNCHED();
    return free_fail(cudaGetLastError(), "ingest_f16 launch");
}

int oevae_quantile_normal(const float* x, float* y, int64_t N, int32_t D, const double* quantiles_t, const double* references, int32_t n_q,
                          double clip_min, double clip_max, int32_t device, void* stream) {
    if (N == 0) return 0;
    if (!x || !y || N < 0 || D <= 0 || !quantiles_t || !references || n_q < 2) { g_create_error = "quantile_normal: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    const long n = (long)N * D;
    { PROF("quantile_normal", (cudaStream_t)stream); scal::quantile_normal_kernel<<<cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, y, n, D, quantiles_t, references, n_q, clip_min, clip_max); }
    LAUNCHED();
    return free_fail(cudaGetLastError(), "quantile_normal launch");
}

int oevae_power_yj(const float* x, float* y, int64_t N, int32_t D, const float* lambdas, const float* mean, const float* scale, int32_t device, void* stream) {
    if (N == 0) return 0;
    if (!x || !y || N < 0 || D <= 0 || !lambdas) { g_create_error = "power_yj: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    const long n = (long)N * D;
    { PROF("power_yj", (cudaStream_t)stream); scal::power_yj_kernel<<<cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, y, n, D, lambdas, mean, scale); }
    LAUNCHED();
    return free_fail(cudaGetLastError(), "power_yj launch");
}

int oevae_pair_metrics(const float* P, const float* Q, int64_t N, int32_t D, float* mse, float* mae, float* mare, float* kld,
                       float* jsd, float* xs, int32_t device, void* stream) {
    if (N == 0) return 0;
    if (!P || !Q || N < 0 || D <= 0) { g_create_error = "pair_metrics: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    MetricOut mo{};
    mo.mse = mse; mo.mae = mae; mo.mare = mare; mo.kld = kld; mo.jsd = jsd; mo.xs = xs;
    score_metrics_kernel<<<cdiv(N * 32, 256), 256, 0, (cudaStream_t)stream>>>(P, Q, 1.f, 0, N, D, mo);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "pair_metrics launch");
}

int oevae_minmax(const float* v, const uint8_t* keep, int64_t N, double* minmax_out, int32_t device, void* stream) {
    if (!minmax_out || (N > 0 && !v)) { g_create_error = "minmax: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    if (N <= 0) return 0;
    unsigned grid = std::min<unsigned>(cdiv(N, 256), 148 * 8);
    minmax_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(v, keep, N, minmax_out);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "minmax launch");
}

int oevae_loss_map(float* v, int64_t N, int32_t branch, int32_t device, void* stream) {
    if ((N > 0 && !v) || branch < 0 || branch > 4) { g_create_error = "loss_map: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    if (N <= 0) return 0;
    loss_map_kernel<<<cdiv(N, 256), 256, 0, (cudaStream_t)stream>>>(v, N, branch);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "loss_map launch");
}

int oevae_hist(const float* var, const float* w, const float* jzw, const float* score, const uint8_t* keep, int64_t N, float cut,
               const double* edges, int32_t n_edges, float wscale, int32_t density, int64_t* counts, double* wsums, int32_t device, void* stream) {
    if (n_edges < 2 || !edges || !counts || (w && !wsums) || (N > 0 && (!var || !jzw))) { g_create_error = "hist: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    if (N <= 0) return 0;
    size_t smem = (size_t)n_edges * 8 + (size_t)(n_edges - 1) * 2 * (8 + 4);
    if (smem > 200 * 1024) { g_create_error = "hist: too many bins"; return OEVAE_EUNSUPPORTED; }
    if (smem > 48 * 1024) { rc = free_fail(cudaFuncSetAttribute(hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "smem attr"); if (rc) return rc; }
    unsigned grid = std::min<unsigned>(cdiv(N, 256 * 8), 148 * 4);
    { PROF("hist", (cudaStream_t)stream); hist_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(var, w, jzw, score, keep, N, cut, edges, n_edges, wscale, density,
                                                           reinterpret_cast<unsigned long long*>(counts), wsums); }
    LAUNCHED();
    return free_fail(cudaGetLastError(), "hist launch");
}

int oevae_wsum(const float* w, const uint8_t* keep, int64_t N, double* sum_acc, int32_t device, void* stream) {
    if (!sum_acc || (N > 0 && !w)) { g_create_error = "wsum: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    if (N <= 0) return 0;
    wsum_kernel<<<std::min<unsigned>(cdiv(N, 1024), 148 * 4), 256, 0, (cudaStream_t)stream>>>(w, keep, N, sum_acc);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "wsum launch");
}

// Layered-GEMM building block exposed for kernel-level tests: C[R,Cn] = sum_t A[r*sAr + t*sAt] * B[t*sBt + c*sBc]
// kind 0/1/2 selects the staging variant used by forward / dX / dW (see Engine::launch_gemm).
int oevae_gemm(oevae_t h, int32_t kind, const float* A, int64_t sAr, int64_t sAt, const float* B, int64_t sBt, int64_t sBc,
               float* C, int32_t ldc, int32_t R, int32_t Cn, int32_t T, const float* bias, int32_t relu, void* stream) {
    H(h);
    if (kind < 0 || kind > 2 || !A || !B || !C || R <= 0 || Cn <= 0 || T < 0) return e->fail(OEVAE_EINVAL, "gemm: bad arguments");
    if ((kind != 2 && sAt != 1) || (kind == 2 && sAr != 1) || (kind == 1 && sBt != 1) || (kind != 1 && sBc != 1))
        return e->fail(OEVAE_EINVAL, "gemm: strides do not match kind %d", kind);
    GemmArgs g{};
    g.A = A; g.B = B; g.C = C; g.sAr = sAr; g.sAt = sAt; g.sBt = sBt; g.sBc = sBc;
    g.R = R; g.Cn = Cn; g.ldc = ldc; g.bias = bias; g.relu = relu; g.t_base = 0; g.t_end = T;
    return e->launch_gemm(g, kind, 1, (cudaStream_t)stream);
}

int oevae_probe_mn(const float* A, const float* B, int32_t K, int32_t M, int32_t N, float* raw_out, int32_t device, void* stream) {
    if (!A || !B || !raw_out || K <= 0 || K > 128 || K % 16 || (M != 64 && M != 128) || N % 16 || N < 16 || N > 128) { g_create_error = "probe: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    int smem = 4 * 128 * 128 + 1024;
    rc = free_fail(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "smem attr"); if (rc) return rc;
    umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, K, M, N, raw_out);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "probe launch");
}

int oevae_probe_latency(int32_t M, int32_t N, int64_t* out_host6, int32_t device) {
    if (!out_host6 || (M != 64 && M != 128) || N % 16 || N < 16 || N > 256) { g_create_error = "probe_latency: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    long long* d = nullptr;
    rc = free_fail(cudaMalloc(&d, 6 * sizeof(long long)), "cudaMalloc"); if (rc) return rc;
    int smem = 2 * 128 * 128 + 1024;
    umma_latency_kernel<<<1, 128, smem>>>(M, N, d);
    LAUNCHED();
    rc = free_fail(cudaMemcpy(out_host6, d, 6 * sizeof(long long), cudaMemcpyDeviceToHost), "probe_latency");
    cudaFree(d);
    return rc;
}

int oevae_probe_issue(int32_t M, int32_t N, int32_t nw, int32_t nm, int32_t mode, int64_t* out_host1, int32_t device) {
    if (!out_host1 || (M != 64 && M != 128) || N % 16 || N < 16 || N > 128 || nw < 1 || nw > 4 || nm < 1) { g_create_error = "probe_issue: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    long long* d = nullptr;
    rc = free_fail(cudaMalloc(&d, sizeof(long long)), "cudaMalloc"); if (rc) return rc;
    umma_issue_kernel<<<1, 128, 2 * 128 * 128 + 1024>>>(M, N, nw, nm, mode, d);
    LAUNCHED();
    rc = free_fail(cudaMemcpy(out_host1, d, sizeof(long long), cudaMemcpyDeviceToHost), "probe_issue");
    cudaFree(d);
    return rc;
}

int oevae_probe_bulk(int32_t bytes, int32_t depth, int32_t n_cta, int32_t same_src, int64_t* out_host2, int32_t device) {
    if (!out_host2 || bytes % 16 || bytes <= 0 || depth < 1 || depth > 8 || (long)bytes * depth > 200 * 1024 || n_cta < 1) { g_create_error = "probe_bulk: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    long long* d = nullptr; unsigned char* src = nullptr;
    size_t total = (size_t)bytes * depth * n_cta;
    rc = free_fail(cudaMalloc(&d, 2 * sizeof(long long)), "cudaMalloc"); if (rc) return rc;
    rc = free_fail(cudaMalloc(&src, total), "cudaMalloc"); if (rc) return rc;
    cudaMemset(src, 1, total);
    int smem = bytes * depth + 1024;
    cudaFuncSetAttribute(bulk_copy_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    bulk_copy_probe_kernel<<<n_cta, 128, smem>>>(src, bytes, depth, same_src ? 0L : (long)bytes * depth, d);
    LAUNCHED();
    rc = free_fail(cudaMemcpy(out_host2, d, 2 * sizeof(long long), cudaMemcpyDeviceToHost), "probe_bulk");
    cudaFree(d); cudaFree(src);
    return rc;
}

// ---- weighted ROC + cut thresholds (roc_kernels.cuh)

int oevae_roc(const float* score, const uint8_t* y_true, int32_t pos_label, const float* w, int64_t N,
              double* fpr, double* tpr, double* thr, int64_t* n_points_host, int32_t device, void* stream) {
    using namespace roc;
    if (!score || !y_true || !fpr || !tpr || !thr || !n_points_host || N <= 0 || N >= (1ll << 31)) { g_create_error = "roc: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const long n = N;
    const unsigned nt = cdiv(n, kTile);
    RocWs ws;
    ws.cap = (size_t)n * (4 * 4 + 2 * 4 + 2 * 8 + 4 + 3 * 8) + (size_t)256 * nt * 4 + (size_t)nt * 24 + (1 << 16);
    rc = free_fail(cudaMalloc(&ws.base, ws.cap), "roc workspace"); if (rc) return rc;
    uint32_t* keys0 = ws.take<uint32_t>(n); uint32_t* keys1 = ws.take<uint32_t>(n);
    uint32_t* vals0 = ws.take<uint32_t>(n); uint32_t* vals1 = ws.take<uint32_t>(n);
    uint32_t* flags = ws.take<uint32_t>(n); uint32_t* pos = ws.take<uint32_t>(n);
    double* a = ws.take<double>(n); double* b = ws.take<double>(n);
    float* ss = ws.take<float>(n);
    double* tps = ws.take<double>(n); double* fps = ws.take<double>(n); double* th = ws.take<double>(n);
    uint32_t* hist = ws.take<uint32_t>((size_t)256 * nt);
    uint32_t* tiles = ws.take<uint32_t>(nt + 64);
    double* dta = ws.take<double>(nt + 1); double* dtb = ws.take<double>(nt + 1);
    uint32_t* total = ws.take<uint32_t>(4);
    int* bad = reinterpret_cast<int*>(ws.take<uint32_t>(4));
    auto done = [&](int code, const char* msg) { cudaStreamSynchronize(s); cudaFree(ws.base); if (code) g_create_error = msg; return code; };
    auto fetch = [&](const void* dptr, void* hptr, size_t bytes) { return cudaMemcpyAsync(hptr, dptr, bytes, cudaMemcpyDeviceToHost, s) == cudaSuccess && cudaStreamSynchronize(s) == cudaSuccess; };
    const unsigned g256 = cdiv(n, 256);
    cudaMemsetAsync(bad, 0, sizeof(int), s);
    // 1. drop zero-weight samples (order-preserving), 2. keys = descending-score order image, payload = sample index
    long m = n;
    if (w) {
        nonzero_flag_kernel<<<g256, 256, 0, s>>>(w, n, flags); LAUNCHED();
        roc_scan(flags, pos, n, tiles, total, s);
        uint32_t h = 0; if (!fetch(total, &h, 4)) return done(OEVAE_ECUDA, "roc: scan failed");
        m = h;
        if (m == 0) return done(OEVAE_EINVAL, "roc: every sample has zero weight");
    }
    build_keys_kernel<<<g256, 256, 0, s>>>(score, n, w ? flags : nullptr, w ? pos : nullptr, keys0, vals0, bad); LAUNCHED();
    const unsigned mt = cdiv(m, kTile);
    uint32_t *kin = keys0, *kout = keys1, *vin = vals0, *vout = vals1;
    for (int shift = 0; shift < 32; shift += 8) {
        rs_hist_kernel<<<mt, kThreads, 0, s>>>(kin, m, shift, hist, (int)mt); LAUNCHED();
        roc_scan(hist, hist, 256L * mt, tiles, nullptr, s);
        rs_scatter_kernel<<<mt, kThreads, 0, s>>>(kin, vin, kout, vout, m, shift, hist, (int)mt); LAUNCHED();
        std::swap(kin, kout); std::swap(vin, vout);
    }
    // 3. class-wise weighted cumulative sums in np.cumsum's order
    gather_sorted_kernel<<<cdiv(m, 256), 256, 0, s>>>(vin, m, score, y_true, pos_label, w, a, b, ss, bad); LAUNCHED();
    int hbad = 0; if (!fetch(bad, &hbad, sizeof(int))) return done(OEVAE_ECUDA, "roc: gather failed");
    if (hbad & 1) return done(OEVAE_EINVAL, "roc: non-finite score (sklearn's roc_curve rejects NaN / inf; apply utils.filtering first)");
    const char* force = getenv("OEVAE_ROC_CUMSUM");                               // "seq" / "par": tests compare the two paths
    const bool seq = force ? (force[0] == 's') : (hbad & 2) != 0;
    if (seq) { seq_cumsum2_kernel<<<1, kThreads, 0, s>>>(a, b, m); LAUNCHED(); }
    else {
        dscan_tiles_kernel<<<mt, kThreads, 0, s>>>(a, b, m, dta, dtb); LAUNCHED();
        dscan_block_kernel<<<1, 32, 0, s>>>(dta, dtb, mt); LAUNCHED();
        dscan_apply_kernel<<<mt, kThreads, 0, s>>>(a, b, m, dta, dtb); LAUNCHED();
    }
    // 4. one point per distinct score
    distinct_flag_kernel<<<cdiv(m, 256), 256, 0, s>>>(ss, m, flags); LAUNCHED();
    roc_scan(flags, pos, m, tiles, total, s);
    uint32_t m1u = 0; if (!fetch(total, &m1u, 4)) return done(OEVAE_ECUDA, "roc: distinct scan failed");
    const long m1 = m1u;
    take_points_kernel<<<cdiv(m, 256), 256, 0, s>>>(flags, pos, m, a, b, ss, tps, fps, th); LAUNCHED();
    // 5. drop_intermediate, 6. normalise
    long m2 = m1;
    if (m1 > 2) {
        corner_flag_kernel<<<cdiv(m1, 256), 256, 0, s>>>(fps, tps, m1, flags); LAUNCHED();
        roc_scan(flags, pos, m1, tiles, total, s);
        uint32_t h = 0; if (!fetch(total, &h, 4)) return done(OEVAE_ECUDA, "roc: corner scan failed");
        m2 = h;
        finish_kernel<<<cdiv(m1, 256), 256, 0, s>>>(flags, pos, m1, fps, tps, th, fpr, tpr, thr); LAUNCHED();
    } else {
        finish_kernel<<<cdiv(m1, 256), 256, 0, s>>>(nullptr, nullptr, m1, fps, tps, th, fpr, tpr, thr); LAUNCHED();
    }
    *n_points_host = m2 + 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return done(OEVAE_ECUDA, cudaGetErrorString(e));
    return done(0, "");
}

// argsort of a float32 vector (ascending, stable) on the device: the m-sort of OoD_pairing (utils.py:195-196)
int oevae_argsort_f32(const float* v, int64_t N, int64_t* order_out, int32_t device, void* stream) {
    using namespace roc;
    if (!v || !order_out || N <= 0 || N >= (1ll << 31)) { g_create_error = "argsort: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const long n = N; const unsigned nt = cdiv(n, kTile);
    RocWs ws; ws.cap = (size_t)n * 16 + (size_t)256 * nt * 4 + (size_t)nt * 8 + (1 << 16);
    rc = free_fail(cudaMalloc(&ws.base, ws.cap), "argsort workspace"); if (rc) return rc;
    uint32_t* k0 = ws.take<uint32_t>(n); uint32_t* k1 = ws.take<uint32_t>(n); uint32_t* v0 = ws.take<uint32_t>(n); uint32_t* v1 = ws.take<uint32_t>(n);
    uint32_t* hist = ws.take<uint32_t>((size_t)256 * nt); uint32_t* tiles = ws.take<uint32_t>(nt + 64);
    asc_keys_kernel<<<cdiv(n, 256), 256, 0, s>>>(v, n, k0, v0); LAUNCHED();
    uint32_t *kin = k0, *kout = k1, *vin = v0, *vout = v1;
    for (int shift = 0; shift < 32; shift += 8) {
        rs_hist_kernel<<<nt, kThreads, 0, s>>>(kin, n, shift, hist, (int)nt); LAUNCHED();
        roc_scan(hist, hist, 256L * nt, tiles, nullptr, s);
        rs_scatter_kernel<<<nt, kThreads, 0, s>>>(kin, vin, kout, vout, n, shift, hist, (int)nt); LAUNCHED();
        std::swap(kin, kout); std::swap(vin, vout);
    }
    widen_idx_kernel<<<cdiv(n, 256), 256, 0, s>>>(vin, n, reinterpret_cast<long long*>(order_out)); LAUNCHED();
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(ws.base);
    return free_fail(e, "argsort");
}

// allocation-free variant of oevae_argsort_f32: workspace = the file-scope g_sort_scratch (see its comment above the C ABI block)
int oevae_argsort_f32_static(const float* v, int64_t N, int64_t* order_out, int32_t device, void* stream) {
    using namespace roc;
    if (!v || !order_out || N <= 0) { g_create_error = "argsort_static: bad arguments"; return OEVAE_EINVAL; }
    if (N > kSortStaticMax) {
        g_create_error = "argsort_static: N = " + std::to_string((long long)N) + " exceeds OEVAE_ARGSORT_STATIC_MAX = " + std::to_string(kSortStaticMax) +
                         " (the static workspace is fixed at build time; use oevae_argsort_f32)";
        return OEVAE_EINVAL;
    }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    RocWs ws; ws.cap = kSortScratchBytes;
    rc = sort_scratch_address(device, &ws.base); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const long n = N; const unsigned nt = cdiv(n, kTile);
    uint32_t* k0 = ws.take<uint32_t>(kSortStaticMax); uint32_t* k1 = ws.take<uint32_t>(kSortStaticMax);
    uint32_t* v0 = ws.take<uint32_t>(kSortStaticMax); uint32_t* v1 = ws.take<uint32_t>(kSortStaticMax);
    uint32_t* hist = ws.take<uint32_t>((size_t)256 * (kSortStaticMax / kTile)); uint32_t* tiles = ws.take<uint32_t>(128);
    if (ws.off > ws.cap) { g_create_error = "argsort_static: workspace layout exceeds the static array"; return OEVAE_EINVAL; }
    asc_keys_kernel<<<cdiv(n, 256), 256, 0, s>>>(v, n, k0, v0); LAUNCHED();
    roc_sort_pairs(k0, v0, k1, v1, n, hist, tiles, s);                               // four passes: the sorted pairs are back in (k0, v0)
    widen_idx_kernel<<<cdiv(n, 256), 256, 0, s>>>(v0, n, reinterpret_cast<long long*>(order_out)); LAUNCHED();
    return free_fail(cudaGetLastError(), "argsort_static launch");                   // asynchronous: the result is ready in stream order
}

int oevae_ood_pairing(const float* m_bkg, const float* pt_bkg, int64_t n_bkg, const float* m_ood_sorted, const float* pt_ood_sorted, int64_t n_ood,
                      uint64_t seed, int32_t pick_first, int64_t* idx_out, float* widths_out, int32_t device, void* stream) {
    if (n_bkg == 0) return 0;
    if (!m_bkg || !pt_bkg || !m_ood_sorted || !pt_ood_sorted || !idx_out || n_bkg < 0 || n_ood <= 0) { g_create_error = "ood_pairing: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    { PROF("ood_pairing", (cudaStream_t)stream);
      pair::ood_pairing_kernel<<<cdiv(n_bkg * 32, 256), 256, 0, (cudaStream_t)stream>>>(m_bkg, pt_bkg, n_bkg, m_ood_sorted, pt_ood_sorted, n_ood, seed, pick_first,
                                                                                       reinterpret_cast<long long*>(idx_out), widths_out); }
    LAUNCHED();
    return free_fail(cudaGetLastError(), "ood_pairing launch");
}

int oevae_digitize_count(const float* var, const uint8_t* keep, int64_t N, const double* edges, int32_t n_edges, uint32_t* counts, int32_t device, void* stream) {
    if (!edges || !counts || n_edges < 2 || N < 0 || (N > 0 && !var)) { g_create_error = "digitize_count: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    if (N == 0) return 0;
    deco::digitize_count_kernel<<<std::min<unsigned>(cdiv(N, 256), 148 * 8), 256, 0, (cudaStream_t)stream>>>(var, keep, N, edges, n_edges, counts);
    LAUNCHED();
    return free_fail(cudaGetLastError(), "digitize_count launch");
}

// mass_deco's device half (plots.py:74-84): cells of all jets, background scores sorted by (cell, score), CDF evaluation in place
int oevae_mass_deco(float* loss, const float* m, const float* pt, const uint8_t* is_bkg, int64_t N, const double* m_edges, int32_t n_m_edges,
                    const double* pt_edges, const int32_t* pt_off, int32_t device, void* stream) {
    using namespace roc;
    if (!loss || !m || !pt || !is_bkg || !m_edges || !pt_edges || !pt_off || N <= 0 || N >= (1ll << 31) || n_m_edges < 2 || n_m_edges > 256) { g_create_error = "mass_deco: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const long n = N; const unsigned nt = cdiv(n, kTile);
    const int n_cells = 256 * 256;
    RocWs ws; ws.cap = (size_t)n * (4 * 8 + 8 + 4) + (size_t)256 * nt * 4 + (size_t)nt * 8 + (size_t)n_cells * 8 + (1 << 16);
    rc = free_fail(cudaMalloc(&ws.base, ws.cap), "mass_deco workspace"); if (rc) return rc;
    uint32_t* cell = ws.take<uint32_t>(n); uint32_t* flags = ws.take<uint32_t>(n); uint32_t* posn = ws.take<uint32_t>(n);
    uint32_t* k0 = ws.take<uint32_t>(n); uint32_t* k1 = ws.take<uint32_t>(n); uint32_t* v0 = ws.take<uint32_t>(n); uint32_t* v1 = ws.take<uint32_t>(n);
    float* sorted = ws.take<float>(n);
    uint32_t* hist = ws.take<uint32_t>((size_t)256 * nt); uint32_t* tiles = ws.take<uint32_t>(nt + 64);
    uint32_t* ccount = ws.take<uint32_t>(n_cells + 1); uint32_t* ctiles = ws.take<uint32_t>(64); uint32_t* total = ws.take<uint32_t>(4);
    auto done = [&](int code, const char* msg) { cudaStreamSynchronize(s); cudaFree(ws.base); if (code) g_create_error = msg; return code; };
    const unsigned g256 = cdiv(n, 256);
    deco::cell_index_kernel<<<g256, 256, 0, s>>>(m, pt, n, m_edges, n_m_edges, pt_edges, pt_off, cell); LAUNCHED();
    // background jets: compact their indices, sort by score, then (stable) by cell
    u8_flag_kernel<<<g256, 256, 0, s>>>(is_bkg, n, flags); LAUNCHED();
    roc_scan(flags, posn, n, tiles, total, s);
    uint32_t nb = 0;
    if (cudaMemcpyAsync(&nb, total, 4, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) return done(OEVAE_ECUDA, "mass_deco: scan failed");
    if (nb == 0) return done(OEVAE_EINVAL, "mass_deco: no background jet");
    compact_score_keys_kernel<<<g256, 256, 0, s>>>(loss, n, flags, posn, k0, v0); LAUNCHED();
    roc_sort_pairs(k0, v0, k1, v1, nb, hist, tiles, s);
    gather_u32_kernel<<<cdiv(nb, 256), 256, 0, s>>>(v0, cell, nb, k0); LAUNCHED();                    // keys = cell of the score-sorted jets
    roc_sort_pairs(k0, v0, k1, v1, nb, hist, tiles, s);
    deco::gather_f32_kernel<<<cdiv(nb, 256), 256, 0, s>>>(v0, loss, nb, sorted); LAUNCHED();
    cudaMemsetAsync(ccount, 0, (n_cells + 1) * sizeof(uint32_t), s);
    deco::cell_count_kernel<<<cdiv(nb, 256), 256, 0, s>>>(k0, nb, ccount); LAUNCHED();
    roc_scan(ccount, ccount, n_cells + 1, ctiles, nullptr, s);                                       // start[c] = #background jets in cells < c
    deco::cdf_apply_kernel<<<g256, 256, 0, s>>>(loss, cell, n, sorted, ccount); LAUNCHED();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return done(OEVAE_ECUDA, cudaGetErrorString(e));
    return done(0, "");
}

int oevae_roc_cuts(const double* fpr, const double* tpr, const double* thr, int64_t n_points, const double* eff_host, int32_t n_eff,
                   double min_tpr, double* out_host, int32_t device, void* stream) {
    if (!fpr || !tpr || !thr || n_points <= 0 || n_eff < 0 || n_eff > 64 || (n_eff && !eff_host) || !out_host) { g_create_error = "roc_cuts: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    double* d = nullptr;
    rc = free_fail(cudaMalloc(&d, (size_t)(n_eff + 4 * (n_eff + 1)) * sizeof(double) + 64), "roc_cuts workspace"); if (rc) return rc;
    double* out = d + n_eff;
    if (n_eff) cudaMemcpyAsync(d, eff_host, n_eff * sizeof(double), cudaMemcpyHostToDevice, s);
    roc::roc_cuts_kernel<<<n_eff + 1, 1024, 0, s>>>(fpr, tpr, thr, n_points, d, n_eff, min_tpr, out); LAUNCHED();
    cudaMemcpyAsync(out_host, out, (size_t)4 * (n_eff + 1) * sizeof(double), cudaMemcpyDeviceToHost, s);
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(d);
    return free_fail(e, "roc_cuts");
}

int oevae_roc_searchsorted(const double* fpr, const double* thr, int64_t n_points, const double* eff_host, int32_t n_q,
                           double* thr_out_host, double* eff_out_host, int32_t device, void* stream) {
    if (!fpr || !thr || n_points <= 0 || n_q <= 0 || n_q > 65536 || !eff_host || !thr_out_host) { g_create_error = "roc_searchsorted: bad arguments"; return OEVAE_EINVAL; }
    int rc = free_fail(cudaSetDevice(device), "cudaSetDevice"); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    double* d = nullptr;
    rc = free_fail(cudaMalloc(&d, (size_t)3 * n_q * sizeof(double)), "roc_searchsorted workspace"); if (rc) return rc;
    cudaMemcpyAsync(d, eff_host, n_q * sizeof(double), cudaMemcpyHostToDevice, s);
    roc::roc_searchsorted_kernel<<<cdiv(n_q, 128), 128, 0, s>>>(fpr, thr, n_points, d, n_q, d + n_q, d + 2 * n_q); LAUNCHED();
    cudaMemcpyAsync(thr_out_host, d + n_q, n_q * sizeof(double), cudaMemcpyDeviceToHost, s);
    if (eff_out_host) cudaMemcpyAsync(eff_out_host, d + 2 * n_q, n_q * sizeof(double), cudaMemcpyDeviceToHost, s);
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(d);
    return free_fail(e, "roc_searchsorted");
}

int64_t oevae_launch_count(void) { return g_launches.load(); }

// page-lock / release a host range in place; "already registered" is success (overlapping loads share pages) and no
// CUDA error is left pending for the caller's next runtime call
int oevae_host_register(void* p, int64_t bytes) {
    if (!p || bytes <= 0) return OEVAE_EINVAL;
    cudaError_t e = cudaHostRegister(p, (size_t)bytes, cudaHostRegisterDefault);
    if (e == cudaSuccess) return 0;
    cudaGetLastError();
    return e == cudaErrorHostMemoryAlreadyRegistered ? 1 : OEVAE_ECUDA;
}
int oevae_host_unregister(void* p) {
    if (!p) return OEVAE_EINVAL;
    cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) { cudaGetLastError(); return OEVAE_ECUDA; }
    return 0;
}

int oevae_profile_enable(int32_t on) {
    if (on && !g_prof.on) g_prof.clear();
    g_prof.on = on != 0;
    return 0;
}

// "name,launches,total_ms\n" per kernel name, sorted by total time; returns the number of bytes written
int oevae_profile_report(char* buf, int32_t cap) {
    if (!buf || cap <= 0) return OEVAE_EINVAL;
    cudaDeviceSynchronize();
    std::map<std::string, std::pair<long, double>> agg;
    for (auto& r : g_prof.recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) != cudaSuccess) continue;
        auto& a = agg[r.name]; a.first++; a.second += ms;
    }
    std::vector<std::pair<std::string, std::pair<long, double>>> v(agg.begin(), agg.end());
    std::sort(v.begin(), v.end(), [](auto& x, auto& y) { return x.second.second > y.second.second; });
    std::string out;
    for (auto& e : v) { char line[256]; snprintf(line, sizeof line, "%s,%ld,%.6f\n", e.first.c_str(), e.second.first, e.second.second); out += line; }
    g_prof.clear();
    int nb = (int)std::min<size_t>(out.size(), (size_t)cap - 1);
    memcpy(buf, out.data(), nb); buf[nb] = 0;
    return nb;
}

}  // extern "C"
