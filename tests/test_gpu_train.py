"""GPU parity tests of the training step (through the C ABI) against the CPU oracle and the golden fixtures.

Tolerances: the fp32 CUDA-core mode computes in the reference's dtype (float32); it is compared with the
float64 oracle at rel 1e-3 / abs floor 1e-6 per jet (BASELINE.json north_star) -- in practice it sits near
1e-5 -- and with looser, separately stated bounds for the flat gradient (cancellation in sums over the batch).
"""
import os
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

from oracle import vae_numpy as V, philox


def _dev(FC, D, oe='KLD', noise='external', precision='fp32', seed=0):
    from atlas_vae_b200.engine import DeviceVAE
    return DeviceVAE(FC, D, oe, precision, noise, seed)


from test_gpu_tensorcore import _close          # records the achieved error in gpurun_out/r02_parity.jsonl


@pytest.mark.parametrize('oe', V.OE_TYPES)
def test_golden_losses_and_gradient(golden_dir, oe):
    g = np.load(os.path.join(golden_dir, 'oracle_train.npz'))
    FC, D = [int(v) for v in g['FC']], int(g['D'])
    dev = _dev(FC, D, oe)
    dev.set_params(g['flat0'])
    B = len(g['X'])
    pj = torch.empty((4, B), device='cuda'); sums = torch.zeros(4, dtype=torch.float64, device='cuda')
    grad = dev.loss_and_grad(g['X'], g['w'], g['Xo'], g['wo'], g['eps'], 0.5, 2.0, 0.3, pj, sums)
    torch.cuda.synchronize()
    _close(pj.cpu().numpy(), g['loss_' + oe], 1e-3, 1e-6, 'per-jet losses')
    _close(pj.cpu().numpy(), g['loss_' + oe], 3e-4, 1e-6, 'per-jet losses (fp32 expectation)')
    _close(sums.cpu().numpy(), g['loss_' + oe].sum(1), 1e-5, 1e-6, 'loss sums')
    gref = g['grad_' + oe]
    scale = np.abs(gref).max()
    _close(grad.cpu().numpy(), gref, 1e-3, 1e-5 * scale, 'flat gradient')
    # eval_losses (no tape) returns the same per-jet numbers
    pj2 = dev.eval_losses(g['X'], g['w'], g['Xo'], g['wo'], g['eps'], 0.5, 2.0, 0.3)
    assert torch.equal(pj2, pj)


def test_golden_adam_trajectory(golden_dir):
    """theta after 1 and 20 Keras-Adam steps and the 20-step loss curve (oracle float64, fixed eps)."""
    g = np.load(os.path.join(golden_dir, 'oracle_train.npz'))
    FC, D = [int(v) for v in g['FC']], int(g['D'])
    dev = _dev(FC, D, 'KLD')
    dev.set_params(g['flat0'])
    B = len(g['X'])
    X, w, Xo, wo, eps = [torch.as_tensor(g[k]).cuda() for k in ('X', 'w', 'Xo', 'wo', 'eps')]
    curve = []
    for s in range(20):
        sums = torch.zeros(4, dtype=torch.float64, device='cuda')
        dev.train_step(X, w, Xo, wo, eps, 1e-3, 0.5, 2.0, 0.3, None, sums)
        curve.append((sums / B).cpu().numpy())
        if s == 0:
            _close(dev.get_params(), g['theta1'], 1e-4, 2e-6, 'theta after 1 step')
    _close(np.array(curve), g['curve20'], 1e-3, 1e-6, '20-step loss curve')
    _close(dev.get_params(), g['theta20'], 1e-3, 2e-4, 'theta after 20 steps')
    m, v, t = dev.get_opt_state()
    assert t == 20 and np.isfinite(m).all() and (v >= 0).all()


@pytest.mark.parametrize('FC,D,B', [([80, 40, 20, 10], 400, 777), ([16], 12, 65), ([33, 7, 5], 23, 1), ([130, 70, 9], 150, 300)])
def test_shapes_vs_oracle(FC, D, B):
    """Ragged / odd shapes incl. B=1, no hidden layer, dims that are not multiples of the GEMM tiles."""
    rng = np.random.default_rng(11)
    flat = V.init_params(FC, D, 3)
    X = rng.normal(0, 0.7, (B, D)).astype(np.float32); X[:, D // 2:] *= (rng.uniform(size=(B, D - D // 2)) > 0.5)
    Xo = rng.normal(0.1, 0.9, (B, D)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, B).astype(np.float32); wo = rng.uniform(0.5, 1.5, B).astype(np.float32)
    eps = rng.normal(size=(B, FC[-1])).astype(np.float32)
    for oe in ('KLD', 'MAE-margin'):
        dev = _dev(FC, D, oe)
        dev.set_params(flat)
        pj = torch.empty((4, B), device='cuda')
        grad = dev.loss_and_grad(X, w, Xo, wo, eps, 0.7, 1.3, 0.2, pj)
        out, gref = V.losses_and_grad(flat, FC, D, X, w, Xo, wo, eps, OE_type=oe, beta=0.7, lamb=1.3, margin=0.2)
        _close(pj.cpu().numpy(), np.stack(out), 1e-3, 1e-6, 'losses %s' % oe)
        _close(grad.cpu().numpy(), gref, 1e-3, 1e-5 * max(np.abs(gref).max(), 1e-3), 'grad %s' % oe)
    # no OoD sample at all: OE term is zero
    dev = _dev(FC, D, 'KLD'); dev.set_params(flat)
    pj = torch.empty((4, B), device='cuda')
    grad = dev.loss_and_grad(X, w, None, None, eps, 0.7, 0., 0., pj)
    out, gref = V.losses_and_grad(flat, FC, D, X, w, None, None, eps, OE_type='KLD', beta=0.7)
    _close(pj.cpu().numpy(), np.stack(out), 1e-3, 1e-6, 'losses no-ood')
    _close(grad.cpu().numpy(), gref, 1e-3, 1e-5 * max(np.abs(gref).max(), 1e-3), 'grad no-ood')


def test_three_pass_rec_oe():
    """MSE outlier exposure with distinct noise for the three vae() calls (reference seed=None behaviour)."""
    rng = np.random.default_rng(5)
    FC, D, B = [24, 12, 6], 20, 97
    flat = V.init_params(FC, D, 1)
    X = rng.normal(size=(B, D)).astype(np.float32); Xo = rng.normal(0.3, 1.2, (B, D)).astype(np.float32)
    w = rng.uniform(.5, 2, B).astype(np.float32); wo = rng.uniform(.5, 2, B).astype(np.float32)
    eps = rng.normal(size=(3, B, 6)).astype(np.float32)
    dev = _dev(FC, D, 'MSE'); dev.set_params(flat)
    pj = torch.empty((4, B), device='cuda')
    grad = dev.loss_and_grad(X, w, Xo, wo, eps, 0.4, 1.7, 0.1, pj)
    out, gref = V.losses_and_grad(flat, FC, D, X, w, Xo, wo, eps[0], eps2=eps[1], eps_ood=eps[2], OE_type='MSE', beta=0.4, lamb=1.7, margin=0.1)
    _close(pj.cpu().numpy(), np.stack(out), 1e-3, 1e-6, 'losses')
    _close(grad.cpu().numpy(), gref, 1e-3, 1e-5 * np.abs(gref).max(), 'grad')


def test_clip_paths():
    """lv = +-200 and huge reconstructions: clip_values forward and its zero gradient (models.py:243-246)."""
    FC, D, B = [8, 4], 6, 16
    flat = V.init_params(FC, D, 0).astype(np.float32)
    lay, _ = V.param_layout(FC, D)
    off = {n: (o, s) for n, s, o in lay}
    flat[off['log_var.bias'][0]:off['log_var.bias'][0] + 4] = [200., -200., 30., 0.]       # exp overflow / underflow / > clip
    flat[off['out.bias'][0]] = 3e6                                                         # clipped reconstruction
    rng = np.random.default_rng(0)
    X = rng.normal(size=(B, D)).astype(np.float32); w = np.ones(B, np.float32)
    eps = rng.normal(size=(B, 4)).astype(np.float32)
    dev = _dev(FC, D, 'KLD'); dev.set_params(flat)
    pj = torch.empty((4, B), device='cuda')
    grad = dev.loss_and_grad(X, w, X[::-1].copy(), w, eps, 1.0, 1.0, 0.5, pj)
    out, gref = V.losses_and_grad(flat, FC, D, X, w, X[::-1].copy(), w, eps, OE_type='KLD', beta=1.0, lamb=1.0, margin=0.5, dtype=np.float32)
    assert np.isfinite(pj.cpu().numpy()).all()
    _close(pj.cpu().numpy(), np.stack(out), 1e-4, 1e-4, 'losses')
    _close(grad.cpu().numpy(), gref, 1e-3, 1e-4 * np.abs(gref).max(), 'grad')
    # NaN gradient entries are sanitised to zero before Adam (models.py:185)
    dev2 = _dev(FC, D, 'KLD'); dev2.set_params(flat)
    gbad = torch.zeros(dev2.n_params, device='cuda'); gbad[0] = float('nan'); gbad[1] = float('inf'); gbad[2] = 5e6; gbad[3] = -2.
    dev2.apply_update(gbad, 1e-3)
    th = dev2.get_params()
    th_ref, _, _ = V.adam_step(flat.astype(np.float64), np.zeros(len(flat)), np.zeros(len(flat)), gbad.cpu().numpy().astype(np.float64), 1, 1e-3)
    _close(th, th_ref, 1e-6, 1e-7, 'adam on sanitised gradient')
    assert th[0] == flat[0] and th[1] == flat[1]


def test_philox_noise_matches_spec():
    """Device Philox/Box-Muller stream vs oracle/philox.py: encoder z = mu + sigma*eps with mu=0, sigma=1."""
    FC, D, B = [4, 10], 3, 1000
    dev = _dev(FC, D, 'KLD', noise='fixed', seed=0x123456789)
    dev.set_params(np.zeros(dev.n_params, np.float32))
    X = np.zeros((B, D), np.float32)
    _, _, z = dev.encode(X)
    e = philox.normal_eps(0x123456789, B, 10)
    np.testing.assert_allclose(z.cpu().numpy(), e, rtol=0, atol=2e-6)
    dev.set_noise_offset(100)
    _, _, z2 = dev.encode(X[:10])
    np.testing.assert_allclose(z2.cpu().numpy(), e[100:110], rtol=0, atol=2e-6)     # sharding invariance
    # per_step mode: consecutive calls differ, fixed mode: identical
    _, _, z3 = dev.encode(X[:10]); assert torch.equal(z2, z3)
    dev_p = _dev(FC, D, 'KLD', noise='per_step', seed=7); dev_p.set_params(np.zeros(dev_p.n_params, np.float32))
    _, _, a = dev_p.encode(X[:10]); _, _, b = dev_p.encode(X[:10])
    assert not torch.equal(a, b)
    np.testing.assert_allclose(b.cpu().numpy(), philox.normal_eps(7, 10, 10, step=4), rtol=0, atol=2e-6)


def test_errors_are_reported():
    from atlas_vae_b200._ffi import OevaeError
    dev = _dev([8, 4], 6, 'KLD')
    with pytest.raises(OevaeError):
        dev.set_params(np.zeros(3, np.float32))
    with pytest.raises(OevaeError):                       # external noise mode without eps
        dev.loss_and_grad(np.zeros((4, 6), np.float32), np.ones(4, np.float32))
    with pytest.raises(OevaeError):                       # x_ood without w_ood
        dev.eval_losses(np.zeros((4, 6), np.float32), np.ones(4, np.float32), np.zeros((4, 6), np.float32), None, np.zeros((4, 4), np.float32))


def test_vae_driver_end_to_end(tmp_path):
    """python -m atlas_vae_b200.vae on BASELINE config 1 shapes (tiny sample): train, score, cuts, histograms."""
    import pickle
    from atlas_vae_b200 import vae as drv
    out = str(tmp_path / 'o')
    rc = drv.main(['--n_train', '4000', '--n_valid', '3000', '--n_sig', '1000', '--n_const', '20', '--n_dims', '4',
                   '--FC_layers', '80', '40', '10', '--batch_size', '1000', '--n_epochs', '2', '--beta', '1', '--apply_cuts', 'ON',
                   '--output_dir', out])
    assert rc == 0
    hist = pickle.load(open(out + '/history.pkl', 'rb'))
    assert len(hist['Train loss']) == 2 and np.isfinite(hist['Train loss']).all() and hist['Train loss'][1] < hist['Train loss'][0]
    res = pickle.load(open(out + '/plots/cut_histograms.pkl', 'rb'))
    assert 'best_gain' in res and res['bkg_eff_1e+01']['n_pass'] > 0
    e, counts, _ = res['bkg_eff_1e+01']['m_uncut']
    assert counts.sum() == 4000                                   # every finite jet lands in a mass bin
