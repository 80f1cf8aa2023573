// Host side of the feature-major fused kernel: plan construction and the weight-image position table.
#pragma once
#include "engine.h"
#include "fused2_plan.h"
#include <algorithm>
#include <vector>

namespace oevae {

static inline uint32_t h2_sw128_off(int r, int c) { return (uint32_t)(r * 128 + ((((c >> 3) ^ (r & 7)) << 4) | ((c & 7) << 1))); }
static inline int h2_r16(int v) { return (v + 15) / 16 * 16; }

// pos: [2][P] byte offsets of the hi / lo bf16 of every kernel parameter inside the image (-1 for biases, which the
// kernel reads in fp32 from the flat vector).  Returns false if the network does not fit the fused kernel.
inline bool build_fused2_plan(const Engine& e, F2Plan& P, std::vector<int>& pos, uint32_t& img_bytes) {
    const int n = e.n, D = e.D, L = e.L;
    if (n < 1 || L > 16 || 2 * n + 2 > kF2MaxStages) return false;
    for (int h : e.hid) if (h > 128) return false;
    P = F2Plan{};
    P.n = n; P.S = 2 * n + 2; P.D = D; P.L = L; P.P = e.P; P.Ppad = e.Ppad;
    P.nX = (D + 63) / 64; P.nC = (D + 127) / 128;
    // ---- stages
    auto seg = [&](F2Stage& s, int k, int lane0, int cnt, const Layer& ly) { s.seg_lane0[k] = lane0; s.seg_n[k] = cnt; s.seg_w[k] = ly.w; s.seg_b[k] = ly.b; };
    auto hidden = [&](int si, int in, int out, const Layer& ly) -> bool {
        F2Stage& s = P.st[si];
        s.in = in; s.out = out; s.kin = h2_r16(in); s.kout = h2_r16(out); s.relu = 1; s.nseg = 1;
        if (std::min(in, out) > 64) return false;
        s.kind = in >= out ? 1 : 2; s.rows = h2_r16(std::max(in, out));
        seg(s, 0, 0, out, ly);
        return true;
    };
    {
        F2Stage& s = P.st[0];
        s.in = D; s.out = e.hid[0]; s.kin = 64 * P.nX; s.kout = h2_r16(s.out); s.kind = 0; s.rows = s.kout; s.relu = 1; s.nseg = 1;
        seg(s, 0, 0, s.out, e.layers[0]);
    }
    for (int k = 1; k < n; ++k) if (!hidden(k, e.hid[k - 1], e.hid[k], e.layers[k])) return false;
    {
        F2Stage& s = P.st[n];
        s.in = e.hid[n - 1]; s.out = L; s.kin = h2_r16(s.in); s.kout = 32; s.kind = 3; s.rows = s.kin; s.relu = 0; s.nseg = 2;
        seg(s, 0, 0, L, e.layers[n]); seg(s, 1, 16, L, e.layers[n + 1]);
    }
    {
        int d = L;
        for (int k = 0; k < n; ++k) {
            int o = e.hid[n - 1 - k];
            if (!hidden(n + 1 + k, d, o, e.layers[n + 2 + k])) return false;
            d = o;
        }
        F2Stage& s = P.st[2 * n + 1];
        s.in = d; s.out = D; s.kin = h2_r16(d); s.kout = 64 * P.nX; s.kind = 4; s.rows = s.kin; s.relu = 0; s.nseg = 1;
        seg(s, 0, 0, D, e.layers[2 * n + 2]);
    }
    // ---- shared memory: tiles, weight ring, S region, scratch
    int off = 0;
    for (int k = 0; k <= 2 * n; ++k) {
        P.tile_rows[k] = (k == n) ? 32 : P.st[k].kout;
        P.tile_off[k] = off; off += 2 * P.tile_rows[k] * 128;
    }
    // ---- ops
    int k = 0; uint32_t img = 1024; uint32_t wmax = 0;                 // image bytes [0,1024) stay zero: the dummy op
    auto add_op = [&](uint32_t bytes) { if (k >= kF2MaxOps) return -1; P.op_off[k] = img; P.op_bytes[k] = bytes; img += bytes; wmax = std::max(wmax, bytes); return k++; };
    auto add_dummy = [&]() { if (k >= kF2MaxOps) return -1; P.op_off[k] = 0; P.op_bytes[k] = 16; return k++; };
    if (2 * P.nX + 4 * n + 8 > kF2MaxOps) return false;
    P.op_first[0] = k;
    for (int c = 0; c < P.nX; ++c) add_op(2u * P.st[0].rows * 128);
    for (int s = 1; s <= 2 * n; ++s) P.op_first[s] = add_op(2u * P.st[s].rows * 128 * (P.st[s].kind == 3 ? 2 : 1));
    if (k & 1) add_dummy();
    P.op_first[2 * n + 1] = k;
    for (int c = 0; c < P.nX; ++c) add_op(2u * P.st[2 * n + 1].rows * 128);
    const int n_fwd = k;
    // backward ops re-read the same image bytes
    for (int s = 2 * n; s >= 1; --s) { P.op_bwd[s] = k; P.op_off[k] = P.op_off[P.op_first[s]]; P.op_bytes[k] = P.op_bytes[P.op_first[s]]; ++k; }
    while (k % kF2WSlots) add_dummy();
    P.n_op_train = k;
    P.n_op_eval = (n_fwd + kF2WSlots - 1) / kF2WSlots * kF2WSlots;   // ops >= n_fwd are shrunk to 16 B by the kernel in eval mode
    img_bytes = img;
    P.wslot_bytes = (int)((wmax + 1023) / 1024 * 1024);
    P.wring_off = off; off += kF2WSlots * P.wslot_bytes;
    P.zero_bytes = P.wring_off;                                        // tiles only: the ring is being filled during set-up
    P.s_off = off; off += std::max(kF2XSlots * 16384, 32768);
    P.scratch_off = off; off += (14080 + 256 * L + 255) / 256 * 256;  // s_w, s_rec, s_kld, s_db [6][4][128] (fused2 uses 4), s_eps, s_wk
    {
        int dwc = 64;
        for (int s = 1; s < P.S; ++s) dwc = std::max(dwc, P.st[s].kin);
        P.dw_cols = dwc; P.n_dwbuf = (2 * dwc <= 192) ? 2 : 1;
        const int stg = 64 * P.st[0].out * 4;
        const int lo = (P.tile_off[1] + 127) / 128 * 128;
        P.stg0_off = (n >= 1 && P.st[0].out % 4 == 0 && lo + 2 * stg <= P.wring_off) ? lo : -1;
    }
    P.smem_bytes = off + 1024;
    if (P.smem_bytes > 227 * 1024) return false;
    // ---- parameter -> image byte positions
    const long Pn = e.P;
    pos.assign(4 * Pn, -1);
    auto put = [&](long param, int op, int block, int nblocks, int rows, int r, int c) {
        uint32_t part = (uint32_t)rows * 128u * nblocks;
        uint32_t hi = P.op_off[op] + (uint32_t)block * rows * 128u + h2_sw128_off(r, c);
        pos[param] = (int)hi; pos[Pn + param] = (int)(hi + part);
    };
    // second copy of a latent column (table entries 2, 3 of the parameter: the update kernels write hi / lo there as well)
    auto put_rep = [&](long param, int op, int block, int nblocks, int rows, int r, int c) {
        uint32_t part = (uint32_t)rows * 128u * nblocks;
        uint32_t hi = P.op_off[op] + (uint32_t)block * rows * 128u + h2_sw128_off(r, c);
        pos[2 * Pn + param] = (int)hi; pos[3 * Pn + param] = (int)(hi + part);
    };
    P.heads_rep = (L <= 16 && P.st[n + 1].kind == 2 && P.st[n + 1].in <= 16) ? 1 : 0;
    for (int s = 0; s < P.S; ++s) {
        const F2Stage& st = P.st[s];
        for (int g = 0; g < st.nseg; ++g) {
            const int out = st.seg_n[g];
            for (int i = 0; i < st.in; ++i)
                for (int j = 0; j < out; ++j) {
                    long prm = st.seg_w[g] + (long)i * out + j;
                    switch (st.kind) {
                        case 0: put(prm, P.op_first[0] + i / 64, 0, 1, st.rows, j, i % 64); break;          // rows = out, K = in
                        case 1: put(prm, P.op_first[s], 0, 1, st.rows, i, j); break;                        // rows = in
                        case 2: put(prm, P.op_first[s], 0, 1, st.rows, j, i);                               // rows = out
                                if (P.heads_rep && s == n + 1) put_rep(prm, P.op_first[s], 0, 1, st.rows, j, i + 16);
                                break;
                        case 3: put(prm, P.op_first[s], g, 2, st.rows, i, j);                               // rows = in, block per head
                                if (P.heads_rep) put_rep(prm, P.op_first[s], g, 2, st.rows, i, j + 16);
                                break;
                        default: put(prm, P.op_first[s] + j / 64, 0, 1, st.rows, i, j % 64); break;         // rows = in, block per 64 outputs
                    }
                }
        }
    }
    return true;
}

}  // namespace oevae
