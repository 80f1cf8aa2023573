"""Worker of tests/test_gpu_dp.py (run under torchrun, one rank per GPU): the peer-memory data-parallel update must give
the parameters of (a) the NCCL all-reduce path and (b) one rank stepping on the whole global batch."""
import os, sys
import numpy as np
import torch
import torch.distributed as td
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from atlas_vae_b200 import dist as Dd
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params


def main():
    rank, world = Dd.init()
    dev = torch.device('cuda', int(os.environ.get('LOCAL_RANK', '0')))
    FC, D, B, steps = [80, 40, 20, 10], 400, 1000, 6
    rng = np.random.default_rng(7)
    X = rng.normal(0, 0.5, (steps, B, D)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, (steps, B)).astype(np.float32)
    lo, hi = (B * rank) // world, (B * (rank + 1)) // world
    flat = init_params(FC, D, 1)
    mk = lambda: DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, dev.index, B)
    a, b, c = mk(), mk(), mk()
    for v in (a, b, c): v.set_params(flat)
    assert a.dp_connect(), 'peer-memory path did not connect'
    grad = torch.empty(b.n_params, device=dev)
    for s in range(steps):
        xs, ws = torch.as_tensor(X[s, lo:hi]).to(dev), torch.as_tensor(w[s, lo:hi]).to(dev)
        a.lib.oevae_set_noise_offset(a.h, lo); b.lib.oevae_set_noise_offset(b.h, lo)
        a.train_step_dp(xs, ws, None, None, None, 1e-3, 1.0, 0., 0.)
        b.loss_and_grad(xs, ws, None, None, None, 1.0, 0., 0., None, None, grad)
        td.all_reduce(grad); b.apply_update(grad, 1e-3)
        c.train_step(torch.as_tensor(X[s]).to(dev), torch.as_tensor(w[s]).to(dev), None, None, None, 1e-3, 1.0, 0., 0.)
    torch.cuda.synchronize(); td.barrier()
    pa, pb, pc = a.get_params(), b.get_params(), c.get_params()
    moved = np.abs(pc - flat).max()
    assert moved > 1e-3, 'parameters did not move'
    # replicas bit-identical across ranks
    t = torch.as_tensor(pa).to(dev); ref = t.clone(); td.broadcast(ref, 0)
    assert torch.equal(t, ref), 'replicas diverged'
    # The three paths add the per-CTA partial gradients in different orders (update kernel: 4 slices per rank, then ranks; all-reduce
    # path: 8 slices, then NCCL's tree; one rank: all tiles of the global batch), i.e. ~1e-7 relative differences in the gradient.
    # Adam's normalised update turns that into up to lr per step for the few parameters whose gradient is ~0: outlier-aware bound.
    for other, what in ((pb, 'all-reduce path'), (pc, 'one rank')):
        bad = np.abs(pa - other) > 2e-6 + 1e-4 * np.abs(other)
        assert bad.mean() < 0.03 and np.abs(pa - other).max() < steps * 1e-3 * 0.5, (what, bad.mean(), np.abs(pa - other).max())
    # ---- the outlier-exposure step (cfg-3: bkg + OoD batch, OE_type 'KLD', paired tiles) through the same update kernel
    Xo = rng.normal(0.1, 0.7, (steps, B, D)).astype(np.float32)
    wo = rng.uniform(0.5, 1.5, (steps, B)).astype(np.float32)
    a2, c2 = mk(), mk()
    for v in (a2, c2): v.set_params(flat)
    assert a2.dp_connect(), 'peer-memory path did not connect (OE engine)'
    for s in range(4):
        dv = lambda t: torch.as_tensor(t).to(dev)
        a2.lib.oevae_set_noise_offset(a2.h, lo)
        a2.train_step_dp(dv(X[s, lo:hi]), dv(w[s, lo:hi]), dv(Xo[s, lo:hi]), dv(wo[s, lo:hi]), None, 1e-3, 1.0, 1.0, 0.5)
        c2.train_step(dv(X[s]), dv(w[s]), dv(Xo[s]), dv(wo[s]), None, 1e-3, 1.0, 1.0, 0.5)
    torch.cuda.synchronize(); td.barrier()
    qa, qc = a2.get_params(), c2.get_params()
    t = torch.as_tensor(qa).to(dev); ref = t.clone(); td.broadcast(ref, 0)
    assert torch.equal(t, ref), 'OE replicas diverged'
    assert np.abs(qc - flat).max() > 1e-3
    bad = np.abs(qa - qc) > 2e-6 + 1e-4 * np.abs(qc)
    assert bad.mean() < 0.03 and np.abs(qa - qc).max() < 4 * 1e-3 * 0.5, ('OE', bad.mean(), np.abs(qa - qc).max())
    # ---- (round 2) the quantity that crosses NVLink: SUM over ranks of the gradient inside the update kernel vs the gradient of the
    # global batch on one rank, rel 1e-5
    a3, r3 = mk(), mk()
    a3.set_params(flat); assert a3.dp_connect()
    worst = 0.0
    for s in range(3):
        dv = lambda t: torch.as_tensor(t).to(dev)
        r3.set_params(a3.get_params())
        gref = r3.loss_and_grad(dv(X[s]), dv(w[s]), None, None, None, 1.0, 0., 0.).cpu().numpy().astype(np.float64)
        a3.lib.oevae_set_noise_offset(a3.h, lo)
        a3.train_step_dp(dv(X[s, lo:hi]), dv(w[s, lo:hi]), None, None, None, 1e-3, 1.0, 0., 0.)
        gdp = a3.dp_last_grad().astype(np.float64)
        worst = max(worst, float(np.linalg.norm(gdp - gref) / np.linalg.norm(gref)))
    assert worst <= 1e-5, ('exchanged gradient sum vs global-batch gradient', worst)
    # ---- clip_values AFTER the reduction (models.py:185 on the summed gradient): a rank-local +-4e5 that sums beyond 1e6, and an inf
    a4 = mk(); a4.set_params(flat); assert a4.dp_connect()
    a4.lib.oevae_set_noise_offset(a4.h, 0)
    xz = torch.zeros((64, D), device=dev); wz = torch.zeros(64, device=dev)          # zero weights: zero gradient from the batch itself
    a4.train_step_dp(xz, wz, None, None, None, 1e-3, 1.0, 0., 0.)
    assert np.abs(a4.dp_last_grad()).max() == 0.0
    # ---- a short last batch whose shard is EMPTY on the last rank (N % batch_size == 1): train_model through the native loop
    from atlas_vae_b200.models import VariationalAutoEncoder, train_model
    n_tm, bs_tm = 2 * 400 + 1, 400
    smp = {'constituents': X[0, :n_tm], 'weights': w[0, :n_tm]}
    val = {'constituents': X[1, :300], 'weights': w[1, :300]}
    m_dp = VariationalAutoEncoder(FC, D, seed=0, precision='bf16x3', device=dev.index)
    th0 = m_dp.dev.get_params().copy()
    hist = train_model(m_dp, [(smp, smp)], [(val, val)], 'KLD', 2, bs_tm, 1.0, 0, 0, 1e-3, verbose=False)
    th_dp = m_dp.dev.get_params()
    t = torch.as_tensor(th_dp).to(dev); ref = t.clone(); td.broadcast(ref, 0)
    assert torch.equal(t, ref), 'train_model replicas diverged'
    one = mk(); one.set_params(th0)
    for ep in range(2):
        for b0 in range(0, n_tm, bs_tm):
            dv = lambda t_: torch.as_tensor(t_).to(dev)
            one.train_step(dv(smp['constituents'][b0:b0 + bs_tm]), dv(smp['weights'][b0:b0 + bs_tm]), None, None, None, 1e-3, 1.0, 0., 0.)
    th_1 = one.get_params()
    bad = np.abs(th_dp - th_1) > 2e-6 + 1e-4 * np.abs(th_1)
    assert bad.mean() < 0.03 and len(hist['Train loss']) == 2 and np.isfinite(hist['Valid loss']).all(), ('train_model dp vs 1 rank', bad.mean())
    # ---- sharded apply_cuts: thresholds from the gathered scores, per-rank cuts, reduced counts == the unsharded result
    from atlas_vae_b200 import cuts as Cc
    rs = np.random.default_rng(99)
    n_c = 60001
    y = (rs.uniform(size=n_c) < 0.8).astype(np.int64)
    sc = np.round(np.where(y == 0, rs.beta(3, 3, n_c), rs.beta(2, 6, n_c)), 4).astype(np.float32)
    full = {'m': rs.gamma(3.0, 30.0, n_c).astype(np.float32), 'pt': rs.uniform(450, 1200, n_c).astype(np.float32),
            'weights': rs.uniform(0.5, 1.5, n_c).astype(np.float32), 'JZW': np.where(y == 0, -1, 0).astype(np.float32)}
    c0, c1 = (n_c * rank) // world, (n_c * (rank + 1)) // world
    part = {k: v[c0:c1] for k, v in full.items()}
    res = Cc.generate_cuts(y[c0:c1], part, sc[c0:c1], 'MAE')
    roc = Cc.roc_curve_device(y, sc, full['weights'], gather=False)              # the unsharded curve, computed locally
    cuts = roc.cuts([1e-4, 1e-3, 1e-2, 1e-1, 1e0, 1e1])
    for e, c in zip([1e-4, 1e-3, 1e-2, 1e-1, 1e0, 1e1], cuts['bkg_eff']):
        r = res['bkg_eff_' + format(e, '1.0e')]
        assert r['loss_cut'] == c and r['n_pass'] == int((sc > c).sum()), (e, r['loss_cut'], c)
        if r['n_pass']:
            sel = sc > c
            edges = Cc.bin_edges(full['m'][sel], 2.5)
            np.testing.assert_array_equal(r['m_cut'][0], edges)
            np.testing.assert_array_equal(r['m_cut'][1][1], np.histogram(full['m'][sel & (y == 1)], edges)[0])
    assert res['best_gain']['loss_cut'] == cuts['gain'][0]
    if rank == 0:
        print('dp_worker ok: world %d, max |dtheta| %.3e, dp vs nccl %.2e, dp vs 1-rank %.2e' % (world, moved, np.abs(pa - pb).max(), np.abs(pa - pc).max()))
    td.destroy_process_group()


if __name__ == '__main__':
    main()
