// Plan of the feature-major fused OE-VAE kernel (fused2_train.cuh), built on the host (fused2_host.cuh).
//
// Orientation: every accumulator has FEATURES on the 128 TMEM lanes and the tile's 64 JETS on the columns
// (D^T = W^T x act^T), so all 128 lanes work in the epilogues, bias / ReLU / bias-gradient are thread-local,
// x is read coalesced and dW^T is stored coalesced.
//
// Shared-memory objects (all bf16 SWIZZLE_128B, hi part then lo part = the bf16x3 operand split):
//   activation tile  [r16(width) feature rows x 64 jets]   MN-major B of the forward / dX products (K = features),
//                                                          K-major A or B of the dW^T products (K = jets)
//   X chunk          [64 jet rows x 64 features]           K-major B of layer 0, MN-major B of dW_0^T
//   weight image     one 64-wide block [rows x 128 B] per op, rows = the larger of (in, out), contiguous = the smaller:
//                    the same bytes are the A operand of the forward product and of the dX product (one K-major,
//                    the other MN-major).  Layer 0: one op per 64 input features (rows = out); output layer: one op
//                    per 64 output features (rows = in); heads: two blocks (mean | log_var).
// Stage s: 0 = first encoder layer, 1..n-1 encoder hidden, n = heads, n+1..2n decoder hidden, 2n+1 = output layer.
// Tile k (k = 0..2n) holds the output of stage k; tile n is the latent tile (z rows 0..15 forward; g_mu rows 0..15,
// g_logvar rows 16..31 backward).
#pragma once
#include <stdint.h>

namespace oevae {

constexpr int kF2MaxStages = 18;
constexpr int kF2MaxOps = 96;
constexpr int kF2Jets = 64;          // jets per tile (UMMA N of the forward products)
constexpr int kF2WSlots = 4;         // weight ring slots
constexpr int kF2XSlots = 2;         // X chunk ring slots

struct F2Stage {
    int in, out;                     // true dims (heads: out = L per head)
    int kin, kout;                   // r16(in); r16(out) (heads: 32)
    int kind;                        // 0 first, 1 hidden rows=in, 2 hidden rows=out, 3 heads, 4 output
    int rows;                        // image rows per block (multiple of 16)
    int relu;
    int nseg;                        // dW^T lane segments -> flat parameters (heads: 2)
    int seg_lane0[2], seg_n[2];
    long seg_w[2], seg_b[2];         // kernel [in, seg_n] and bias [seg_n] offsets in the flat vector
};

struct F2Plan {
    int n, S, D, L;                  // hidden layers, stages (2n+2), input dim, latent dim
    int nX;                          // 64-feature chunks of X (= 64-feature blocks of the output layer)
    int nC;                          // 128-feature chunks of the output layer
    int tile_off[kF2MaxStages], tile_rows[kF2MaxStages];
    int wring_off, wslot_bytes;      // kF2WSlots slots
    int s_off;                       // X ring (kF2XSlots x 16 KB) aliased with the g_y slot (32 KB)
    int scratch_off;                 // fp32 scratch (see kernel)
    int zero_bytes;                  // [0, zero_bytes): cleared once per CTA
    int smem_bytes;
    int n_op_train, n_op_eval;       // both multiples of kF2WSlots; op k lives in ring slot k % kF2WSlots
    int op_first[kF2MaxStages];      // per stage: first forward op (stage 0: nX ops; output: nX ops)
    int op_bwd[kF2MaxStages];        // per stage: backward op (hidden + heads)
    uint32_t op_off[kF2MaxOps], op_bytes[kF2MaxOps];
    F2Stage st[kF2MaxStages];
    long P, Ppad;
    int n_dwbuf, dw_cols;            // fused3: dW^T accumulators in TMEM columns [320, 320 + n_dwbuf * dw_cols)
    int heads_rep;                   // fused3: the heads' and stage n+1's weight images carry a second copy of their latent columns (16..31),
                                     // so accumulator rows 16 + l replicate rows l and the Sampling / latent-backward epilogues use 32 lanes
    int stg0_off;                    // fused3: two [64 x out_0] fp32 staging buffers for the layer-0 gradient chunks inside the
                                     // tiles 1..2n (dead by then), or -1 when they do not fit / out_0 % 4 != 0 (direct stores)
};

}  // namespace oevae
