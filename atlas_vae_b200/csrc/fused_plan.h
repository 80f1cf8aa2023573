// Static schedule ("plan") of the fused per-tile OE-VAE training kernel (fused_train.cuh), built on the host.
//
// Fused layer view: enc hidden 0..n-1, heads (dense_mean | dense_log_var merged, log_var columns start at Lp =
// roundup16(L)), dec hidden 0..n-1, output.  nl = 2n + 2 fused layers.
// Activation tensors ("tiles") live in shared memory as [Rt rows x 64-column blocks] bf16 hi/lo in the canonical
// SWIZZLE_128B layout; tile k holds the OUTPUT of fused layer k (k = 0..nl-2) plus a ones column at index `width`
// (bias via the contraction).  The heads' output tile holds z (L columns + ones) and later g_heads (2*Lp columns).
#pragma once
#include <stdint.h>

namespace oevae {

constexpr int kMaxFL = 18;       // fused layers
constexpr int kMaxWop = 128;     // weight-blob operations per tile

struct FLayerDev {
    int in, out;                 // true dims (heads: out = Lp + L, i.e. last real column + 1)
    int npad, inpad;             // UMMA N of the forward / dX contractions (multiples of 16)
    int kb_fwd, kb_dx;           // 64-wide K blocks of the forward (in+1) and dX (out) contractions
    int relu;
    // dW accumulator columns -> flat parameter offsets (heads have two segments: mean, log_var)
    int seg_col0[2], seg_n[2], seg_ld[2];
    long seg_dst[2];
};

struct FPlan {
    int nl, n_hidden, D, L, Lp, Rt, blk;        // blk = Rt*128 bytes (one 64-column block)
    int nX, nY;                                  // 64-column chunks of [X,1] (D+1) and of the output (D)
    int tile_off[kMaxFL], tile_nblk[kMaxFL];     // smem byte offset of the hi part; lo part follows at +nblk*blk
    int wide_off;                                // 2 slots x (hi blk, lo blk)
    int wring_off, wslot_bytes;                  // 2 weight slots
    int scratch_off;                             // fp32 scratch: mu, lv, eps [Rt][L] each
    int smem_bytes;
    int n_wop;                                   // weight ops per tile (train mode order, see fused_train.cuh)
    uint32_t wop_off[kMaxWop], wop_bytes[kMaxWop];
    FLayerDev layer[kMaxFL];
    long P, Ppad;
};

}  // namespace oevae
