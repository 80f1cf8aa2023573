// Host side of the fused training kernel: plan construction, weight-image position tables, launches.
#pragma once
#include "engine.h"
#include "fused_train.cuh"
#include <algorithm>
#include <vector>

namespace oevae {

static inline uint32_t h_sw128_off(int r, int c) { return (uint32_t)(r * 128 + ((((c >> 3) ^ (r & 7)) << 4) | ((c & 7) << 1))); }
static inline int r16(int v) { return (v + 15) / 16 * 16; }

// Builds the plan for tile height Rt; returns false if the network does not fit the fused kernel.
inline bool build_fused_plan(const Engine& e, int Rt, FPlan& P, std::vector<int>& pos) {
    const int n = e.n, D = e.D, L = e.L;
    if (n < 1 || L > 32) return false;
    for (int h : e.hid) if (h > 127) return false;
    P = FPlan{};
    P.n_hidden = n; P.nl = 2 * n + 2; P.D = D; P.L = L; P.Lp = r16(L); P.Rt = Rt; P.blk = Rt * 128;
    P.nX = (D + 1 + 63) / 64; P.nY = (D + 63) / 64;
    P.P = e.P; P.Ppad = e.Ppad;
    if (P.nl > kMaxFL) return false;
    // fused layers
    auto setl = [&](int f, int in, int out, int relu) {
        FLayerDev& l = P.layer[f];
        l.in = in; l.out = out; l.npad = r16(out); l.inpad = r16(in);
        l.kb_fwd = (in + 1 + 63) / 64; l.kb_dx = (out + 63) / 64; l.relu = relu;
    };
    int d = D;
    for (int k = 0; k < n; ++k) {
        setl(k, d, e.hid[k], 1);
        P.layer[k].seg_col0[0] = 0; P.layer[k].seg_n[0] = e.hid[k]; P.layer[k].seg_ld[0] = e.hid[k]; P.layer[k].seg_dst[0] = e.layers[k].w;
        d = e.hid[k];
    }
    setl(n, d, P.Lp + L, 0);
    P.layer[n].seg_col0[0] = 0; P.layer[n].seg_n[0] = L; P.layer[n].seg_ld[0] = L; P.layer[n].seg_dst[0] = e.layers[n].w;
    P.layer[n].seg_col0[1] = P.Lp; P.layer[n].seg_n[1] = L; P.layer[n].seg_ld[1] = L; P.layer[n].seg_dst[1] = e.layers[n + 1].w;
    d = L;
    for (int k = 0; k < n; ++k) {
        int f = n + 1 + k, li = n + 2 + k, o = e.hid[n - 1 - k];
        setl(f, d, o, 1);
        P.layer[f].seg_col0[0] = 0; P.layer[f].seg_n[0] = o; P.layer[f].seg_ld[0] = o; P.layer[f].seg_dst[0] = e.layers[li].w;
        d = o;
    }
    {
        int f = 2 * n + 1, li = 2 * n + 2;
        setl(f, d, D, 0);
        P.layer[f].seg_col0[0] = 0; P.layer[f].seg_n[0] = D; P.layer[f].seg_ld[0] = D; P.layer[f].seg_dst[0] = e.layers[li].w;
    }
    for (int f = 0; f < P.nl - 1; ++f) if (P.layer[f].npad > 128 || P.layer[f].inpad > 128 && f > 0) return false;
    // shared memory carve-up
    int off = 0;
    for (int k = 0; k <= P.nl - 2; ++k) {
        int width = (k == n) ? std::max(L + 1, 2 * P.Lp) : P.layer[k].out + 1;
        P.tile_nblk[k] = (width + 63) / 64;
        if (P.tile_nblk[k] > 2) return false;
        P.tile_off[k] = off; off += 2 * P.tile_nblk[k] * P.blk;
    }
    P.wide_off = off; off += 2 * 2 * P.blk;
    off = (off + 1023) / 1024 * 1024;
    P.wring_off = off;
    // weight ops (train order) and blob layout
    uint32_t img = 0; int k = 0; uint32_t wmax = 0;
    auto add_op = [&](uint32_t bytes) { P.wop_off[k] = img; P.wop_bytes[k] = bytes; img += bytes; wmax = std::max(wmax, bytes); ++k; };
    if (P.nX + 2 * (P.nl - 2) + P.nY > kMaxWop) return false;
    std::vector<uint32_t> op_f0(P.nX), op_fwd(P.nl), op_out(P.nY), op_dx(P.nl);
    for (int c = 0; c < P.nX; ++c) { op_f0[c] = img; add_op(2u * P.layer[0].npad * 128); }
    for (int f = 1; f <= P.nl - 2; ++f) { op_fwd[f] = img; add_op(2u * P.layer[f].kb_fwd * P.layer[f].npad * 128); }
    const FLayerDev& lo = P.layer[P.nl - 1];
    for (int c = 0; c < P.nY; ++c) { op_out[c] = img; add_op(2u * lo.kb_fwd * 64 * 128 + 2u * lo.inpad * 128); }
    for (int f = P.nl - 2; f >= 1; --f) { op_dx[f] = img; add_op(2u * P.layer[f].kb_dx * P.layer[f].inpad * 128); }
    P.n_wop = k;
    P.wslot_bytes = (int)((wmax + 1023) / 1024 * 1024);
    off += 2 * P.wslot_bytes;
    P.scratch_off = off; off += (3 * Rt * L + 2 * Rt) * 4;
    P.smem_bytes = off + 1024;
    if (P.smem_bytes > 227 * 1024) return false;
    // ---- per-parameter byte positions inside the image: [fwd hi, fwd lo, dx hi, dx lo]
    const long Pn = e.P;
    pos.assign(4 * Pn, -1);
    auto put_fwd = [&](int f, long param, int i /*K row: input feature or 'in' for bias*/, int j /*fused output column*/) {
        const FLayerDev& l = P.layer[f];
        uint32_t hi, lo_;
        if (f == 0) { uint32_t b = op_f0[i / 64]; hi = b + h_sw128_off(j, i % 64); lo_ = hi + l.npad * 128; }
        else if (f <= P.nl - 2) { uint32_t b = op_fwd[f] + (i / 64) * l.npad * 128; hi = b + h_sw128_off(j, i % 64); lo_ = hi + l.kb_fwd * l.npad * 128; }
        else { uint32_t b = op_out[j / 64] + (i / 64) * 64 * 128; hi = b + h_sw128_off(j % 64, i % 64); lo_ = hi + l.kb_fwd * 64 * 128; }
        pos[param] = (int)hi; pos[Pn + param] = (int)lo_;
    };
    auto put_dx = [&](int f, long param, int i, int j) {
        const FLayerDev& l = P.layer[f];
        uint32_t hi, lo_;
        if (f == 0) return;
        if (f <= P.nl - 2) { uint32_t b = op_dx[f] + (j / 64) * l.inpad * 128; hi = b + h_sw128_off(i, j % 64); lo_ = hi + l.kb_dx * l.inpad * 128; }
        else { uint32_t b = op_out[j / 64] + 2u * l.kb_fwd * 64 * 128; hi = b + h_sw128_off(i, j % 64); lo_ = hi + l.inpad * 128; }
        pos[2 * Pn + param] = (int)hi; pos[3 * Pn + param] = (int)lo_;
    };
    auto map_layer = [&](int f, const Layer& ly, int col_shift) {
        for (int i = 0; i < ly.in; ++i)
            for (int j = 0; j < ly.out; ++j) { put_fwd(f, ly.w + (long)i * ly.out + j, i, j + col_shift); put_dx(f, ly.w + (long)i * ly.out + j, i, j + col_shift); }
        for (int j = 0; j < ly.out; ++j) put_fwd(f, ly.b + j, ly.in, j + col_shift);
    };
    for (int kk = 0; kk < n; ++kk) map_layer(kk, e.layers[kk], 0);
    map_layer(n, e.layers[n], 0);
    map_layer(n, e.layers[n + 1], P.Lp);
    for (int kk = 0; kk < n; ++kk) map_layer(n + 1 + kk, e.layers[n + 2 + kk], 0);
    map_layer(2 * n + 1, e.layers[2 * n + 2], 0);
    return true;
}

}  // namespace oevae
