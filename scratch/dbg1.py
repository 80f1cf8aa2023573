import sys, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from oracle import vae_numpy as V
rng = np.random.default_rng(0)
print('--- gemm unit')
for prec in ('bf16x3', 'bf16', 'fp32'):
    dev = DeviceVAE([8, 4], 6, 'KLD', prec, 'external', 0)
    for (R, T, Cn) in [(5, 7, 3), (128, 64, 16), (300, 400, 80), (2100, 256, 400), (401, 2100, 256), (129, 1, 5), (64, 200, 130)]:
        for kind in (0, 1, 2):
            A = rng.normal(size=(R, T)); Bm = rng.normal(size=(T, Cn))
            ref = A @ Bm
            if kind == 0: a, b = A, Bm
            elif kind == 1: a, b = A, Bm.T.copy()
            else: a, b = A.T.copy(), Bm
            out = dev.gemm(kind, a.astype(np.float32), b.astype(np.float32)).cpu().numpy()
            err = np.abs(out - ref).max() / np.abs(ref).max()
            flag = '' if err < (1e-4 if prec != 'bf16' else 2e-2) else '   <<<<<< BAD'
            print(prec, kind, (R, T, Cn), 'rel-to-max err %.2e' % err, flag)
print('--- failing shapes: per-layer gradient error, bf16x3 vs fp32 vs oracle')
for FC, D, B in [([33, 7, 5], 23, 1), ([256, 128, 64], 400, 2100)]:
    rng = np.random.default_rng(11)
    flat = V.init_params(FC, D, 3)
    X = rng.normal(0, 0.7, (B, D)).astype(np.float32); X[:, D // 2:] *= (rng.uniform(size=(B, D - D // 2)) > 0.5)
    Xo = rng.normal(0.1, 0.9, (B, D)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, B).astype(np.float32); wo = rng.uniform(0.5, 1.5, B).astype(np.float32)
    eps = rng.normal(size=(B, FC[-1])).astype(np.float32)
    for margin in (0.2, 50.0):
        out, gref = V.losses_and_grad(flat, FC, D, X, w, Xo, wo, eps, OE_type='KLD', beta=0.7, lamb=1.3, margin=margin)
        res = {}
        for prec in ('fp32', 'bf16x3'):
            dev = DeviceVAE(FC, D, 'KLD', prec, 'external', 0); dev.set_params(flat)
            pj = torch.empty((4, B), device='cuda')
            res[prec] = dev.loss_and_grad(X, w, Xo, wo, eps, 0.7, 1.3, margin, pj).cpu().numpy()
            lay, _ = V.param_layout(FC, D)
            print(FC, B, 'margin', margin, prec, 'max|g|=%.3e' % np.abs(gref).max(), 'loss err %.2e' % np.abs(pj.cpu().numpy() - np.stack(out)).max())
            for name, shape, off in lay:
                n = int(np.prod(shape)); e = np.abs(res[prec][off:off + n] - gref[off:off + n]).max(); sc = np.abs(gref[off:off + n]).max()
                print('    %-16s err %.3e  scale %.3e' % (name, e, sc))
