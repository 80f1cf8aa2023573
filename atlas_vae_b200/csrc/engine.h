// Host-side state behind an oevae_t handle.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <string>
#include <vector>
#include "../../include/oevae.h"

namespace oevae {

struct GemmArgs;
struct FusedState;
struct EarlyGroups;
constexpr int kMaxSlots = 48;          // split-K partial-gradient slots (3 passes x 16)

struct Layer { int in, out; long w, b; };   // offsets of kernel [in,out] and bias [out] in the flat vector

// one call = up to three row groups of B jets each (SURVEY 7.1 / ref_kernels.cuh combine_kernel)
struct Pass {
    long B = 0;
    int R = 1;                // row groups that run the encoder
    int Rd = 1;               // leading groups that also run the decoder
    const float* x[3] = {nullptr, nullptr, nullptr};
    const float* eps = nullptr; int n_eps = 0;
    bool want_rec = true;
    bool encoder_only = false;
};

// Tuning / debugging switches, read from the environment ONCE when the handle is created (never on the launch path):
//   OEVAE_FUSED=0          layered path even where the fused kernel applies      OEVAE_FUSED_V=2       fused2 (cross-check) instead of fused3
//   OEVAE_OE_LAYERED=1     outlier-exposure step on the layered path             OEVAE_SCORE_LAYERED=1 scoring on the layered path
//   OEVAE_SCORE_STATIC=1   round-robin tiles in the scoring pass                 OEVAE_SCORE_HEADS_T=1 scoring heads with the jets on the lanes
//   OEVAE_ADAM_EARLY=0     every block of the update kernel waits for the whole training grid (A/B of the early parameter groups)
//   OEVAE_FUSED_DBG=1      phase stamps of CTA 0 to stderr (synchronises)        OEVAE_DBG_SKIP=k      stamps from the k-th tile of CTA 0 on
struct Knobs {
    bool fused = true, oe_layered = false, score_layered = false, score_static = false, fused_dbg = false, dp_stamps = false;
    int adam_early = 27;      // bit 0: output-layer group, bit 1: chain group, bit 2: dry run (poll only), bit 3: groups also in the data-parallel
                              // update, bit 4: finished-CTA counter instead of griddepcontrol.wait (both updates)
    int fused_v = 3, score_heads_t = 0;
    long long dbg_skip = 0;
    void read_env();
};

struct Engine {
    oevae_config cfg{};
    Knobs knobs;
    std::string err;
    int D = 0, L = 0, n = 0, maxH = 0, n_sm = 0;
    std::vector<int> hid;
    std::vector<Layer> layers;
    long P = 0, Ppad = 0;
    float *theta = nullptr, *m = nullptr, *v = nullptr;
    long long step = 0;           // Adam t
    long long calls = 0;          // forward-call counter (Philox step in PER_STEP mode)
    unsigned long long row_offset = 0;

    // workspace
    float* arena = nullptr; long cap_rows = 0;
    float* parts = nullptr;
    std::vector<float*> act_e, act_d;
    float *mu = nullptr, *lv = nullptr, *z = nullptr, *eps_buf = nullptr, *g_mu = nullptr, *g_lv = nullptr;
    float *y = nullptr, *ysum = nullptr, *gA = nullptr, *gB = nullptr;
    float *kld = nullptr, *rec = nullptr, *g_rec = nullptr, *g_kld = nullptr;
    int Sg = 1, chunk_rows = 0, n_slots_used = 0;
    FusedState* fz = nullptr;      // fused tcgen05 training kernel (bf16x3, small nets); null when not applicable
    // data-parallel exchange over NVLink peer memory (dp_kernels.cuh): receive buffers [2 parities][8 ranks][slots] {three sums, tag}
    float* dp_buf = nullptr; void* dp_peer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int dp_rank = -1, dp_world = 0, dp_nblk = 0; unsigned int dp_seq = 0; unsigned long long dp_count = 0;
    long long* dp_stamps = nullptr;   // OEVAE_DP_STAMPS=1: per-block phase stamps of the last update kernel (oevae_dp_stamps_host)
    float* dp_gsum = nullptr;     // [P] SUM over ranks of the last step's gradient, before clip_values (parity checks: oevae_dp_last_grad_host)

    // host-load loop (oevae_train_batches_host): copy stream, per-slot events, staging ring of kHlSlots batches; the copies run
    // kHlSlots - 1 batches ahead of the steps, so the copy engine stays busy while the host reads the running means (a sync every
    // 10 batches, models.py:191)
    static constexpr int kHlSlots = 6;
    cudaStream_t hl_copy = nullptr, hl_prep = nullptr; cudaEvent_t hl_ready[kHlSlots] = {}, hl_freed[kHlSlots] = {}, hl_copied[kHlSlots] = {};
    const double* hl_trace_src = nullptr; double* hl_trace_dst = nullptr;   // set around the update launch of a host-load step (adam_kernels.cuh SumsTrace)
    unsigned char* hl_ring = nullptr; size_t hl_ring_bytes = 0;
    bool hl_used[kHlSlots] = {};
    const void* hl_src = nullptr; long hl_copied_upto = 0;   // batches [.., hl_copied_upto) of the load at hl_src have their copies issued

    ~Engine();
    int fail(int code, const char* fmt, ...);
    int cuda_check(cudaError_t e, const char* what);
    int init(const oevae_config& c);
    int reserve(long rows);
    int mse_type() const { return cfg.oe_type == OEVAE_OE_MSE || cfg.oe_type == OEVAE_OE_MSE_MARGIN; }

    int launch_gemm(const GemmArgs& g, int kind, int n_slots, cudaStream_t s);
    int gemm_fwd(const float* A, int li, float* C, long rows, bool relu, cudaStream_t s);
    int gemm_dx(const float* G, int li, float* C, long rows, bool accumulate, const float* mask, cudaStream_t s);
    int gemm_dw(const float* A, const float* G, int li, long rows, long B, int slot0, int n_slots, cudaStream_t s);

    int plan(Pass& p, const float* x_bkg, const float* x_ood, const float* eps, int n_eps, long B);
    int forward(const Pass& p, cudaStream_t s);
    int forward_encoder(const Pass& p, cudaStream_t s);
    int backward(const Pass& p, cudaStream_t s);
    int losses(const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood, const float* eps, int n_eps,
               long B, float beta, float lamb, float margin, float* per_jet, double* sums, bool want_grad, cudaStream_t s);
    int fused_init();
    int refresh_images(cudaStream_t s);
    long long* dbg_buffer();
    int fused_losses(const float* x, const float* w, const float* eps, long B, float beta, float* per_jet, double* sums,
                     bool want_grad, cudaStream_t s);
    int fused_oe(const float* x_bkg, const float* w_bkg, const float* x_ood, const float* w_ood, const float* eps, long B,
                 float beta, float lamb, float margin, float* per_jet, double* sums, bool want_grad, cudaStream_t s);
    EarlyGroups early_groups() const;
    int reduce_and_update(float* grad_out, const float* grad_in, bool apply, float lr, cudaStream_t s);
    int dp_export(void* handle_out64);
    int dp_connect(int rank, int world, const void* handles);
    int dp_update(float lr, bool have_grad, cudaStream_t s);
    int reset_opt();
    int train_batches_host(oevae_host_load& a, cudaStream_t s);
    int score(const float* x, long N, int n_iter, const float* eps, float* o_mse, float* o_mae, float* o_kld, float* o_jsd,
              float* o_lat, unsigned char* o_fin, float* o_xhat, cudaStream_t s);
    int encode(const float* x, long N, const float* eps, float* o_mu, float* o_lv, float* o_z, cudaStream_t s);
};

}  // namespace oevae
