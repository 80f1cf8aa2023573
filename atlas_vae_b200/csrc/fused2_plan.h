// Plan of the feature-major fused training kernel (fused2_train.cuh).
//
// Orientation: every accumulator has FEATURES on the TMEM lanes and JETS on the columns (D = W^T-image x act), so
// that all 128 lanes work in the epilogues, activation rows are written with 16-byte stores, x is read coalesced
// and dW^T rows are stored coalesced.  Activation "T-tiles" live in shared memory as [feature rows x 64 jets] bf16
// SWIZZLE_128B (hi part, then lo part); row `width` of a tile is all ones (bias gradient through the dW product).
// Keras layer index li: enc k (0..n-1), mean (n), log_var (n+1), dec k (n+2..2n+1), output (2n+2).
#pragma once
#include <stdint.h>

namespace oevae {

constexpr int kF2MaxLayers = 19;
constexpr int kF2MaxOps = 96;

struct F2Layer {
    int in, out;
    int M, Min;                  // lanes (64/128) of the forward accumulator (>= out) and of the dX accumulator (>= in)
    int kb_in, kb_out;           // ceil(in/64), ceil(out/64)
    int in_tile, out_tile;       // T-tile ids; in_tile -1 = X (layer 0); out_tile -1 = output layer
    int relu;
    int npad_in;                 // r16(in + 1): UMMA N of the dW^T product
    long w_off, b_off;
};

struct F2Plan {
    int n, nlay, D, L;
    int nX;                      // 64-feature chunks of X
    int nC;                      // 128-column chunks of the output layer
    int n_tiles_smem;            // number of T-tiles
    int tile_off[kF2MaxLayers];  // byte offset of the hi part
    int tile_lo[kF2MaxLayers];   // byte distance hi -> lo part (= rows_alloc * 128)
    int wide_off;                // X ring (2 x [64 jets x 128 B] hi+lo) aliased with the g_y slot ([128 x 128 B] hi+lo)
    int wring_off;               // 2 weight slots of 32 KB
    int scratch_off;             // fp32: mu, lv, eps [L][64]; w [64]; kld [64]; rec [4][64]
    int zero_bytes;              // [0, zero_bytes) is cleared once per CTA (tiles + wide region)
    int smem_bytes;
    int n_op_train, n_op_eval;
    uint32_t op_off_train[kF2MaxOps], op_bytes_train[kF2MaxOps];
    uint32_t op_off_eval[kF2MaxOps], op_bytes_eval[kF2MaxOps];
    F2Layer layer[kF2MaxLayers];
    long P, Ppad;
};

}  // namespace oevae
