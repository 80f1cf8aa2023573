"""CPU tests: the oracle against the reference-generated golden vectors (tests/golden/make_golden.py),
the Random123 Philox KATs, and the two independent gradient derivations against each other."""
import json, os
import numpy as np
import pytest

from oracle import vae_numpy as V, philox, scoring as S, preprocess as PP


def test_philox_random123_kat():
    # Random123 kat_vectors, philox4x32-10
    kat = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
            [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for c, k, exp in kat:
        got = philox.philox4x32_10(np.array(c, np.uint32), np.array(k, np.uint32))
        assert [int(v) for v in got] == exp


def test_philox_normals_are_standard():
    e = philox.normal_eps(0, 50000, 10)
    assert abs(e.mean()) < 0.01 and abs(e.std() - 1) < 0.01
    assert np.array_equal(e[100:110], philox.normal_eps(0, 10, 10, row_offset=100))      # sharding invariance
    assert not np.array_equal(e[:10], philox.normal_eps(0, 10, 10, step=1))


def test_scoring_metrics_match_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    got = S.score_metrics(g['P'], g['Q'], ['MSE', 'MAE', 'MARE', 'KLD', 'JSD', 'X-S'])
    for m, v in got.items():
        np.testing.assert_array_equal(v, g['m_' + m], err_msg=m)
        assert v.dtype == g['m_' + m].dtype
    keep = S.finite_rows(g['filt_bad'])
    np.testing.assert_array_equal(np.arange(len(keep))[keep], g['filt_kept'])
    for i in range(5):
        np.testing.assert_array_equal(S.loss_mapping(g['map_in%d' % i]), g['map_out%d' % i])
        assert S.loss_mapping_branch(g['map_in%d' % i]) == i


def test_roc_cut_hist_match_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    rates = S.get_rates(g['roc_y'], g['roc_score'], g['roc_w'])
    for a, b in zip(rates, (g['roc_fpr'], g['roc_tpr'], g['roc_thr'])):
        np.testing.assert_array_equal(a, b)
    thr, val = S.best_threshold(rates, 'gain')
    assert thr == g['thr_gain'] and val == g['val_gain']
    idx = np.arange(len(g['roc_y']))
    np.testing.assert_array_equal(idx[S.cut_mask(g['roc_score'], S.cut_threshold(rates))], g['kept_gain'])
    for eff in [1e-1, 1e0, 1e1]:
        m = S.cut_mask(g['roc_score'], S.cut_threshold(rates, eff))
        np.testing.assert_array_equal(idx[m], g['kept_eff_%g' % eff])
    for tag, bs in [('m', 2.5), ('pt', 10)]:
        e = S.bin_edges(g['var_' + tag], bs)
        assert e.dtype == np.float64
        np.testing.assert_array_equal(e, g['edges_' + tag])
        JZW = np.zeros(len(g['var_' + tag]), np.float32)
        c, ws = S.class_histograms(g['var_' + tag], np.ones_like(JZW), JZW, e)
        np.testing.assert_array_equal(c[1], g['hist_' + tag]); assert c[0].sum() == 0
        np.testing.assert_array_equal(ws[1], c[1].astype(np.float64))


def test_preprocess_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, 'ref_preprocess.npz'))
    np.testing.assert_array_equal(PP.sort_constituents(g['raw']), g['sorted'])
    np.testing.assert_array_equal(PP.sort_constituents(g['raw16']), g['sorted16'])
    x, pt, m = PP.preprocess(g['raw'], 20, 4)
    np.testing.assert_array_equal(x, g['sorted'][:, :80])
    np.testing.assert_array_equal(pt, g['pt20']); np.testing.assert_array_equal(m, g['m20'])
    x3, _, _ = PP.preprocess(g['raw'], 20, 3)
    np.testing.assert_array_equal(x3, g['sorted'][:, :80].reshape(-1, 20, 4)[..., 1:].reshape(len(x), -1))
    xp, _, _ = PP.preprocess(g['raw'], 60, 4)            # n_const beyond n_raw -> zero padding (utils.py:93-96)
    assert xp.shape == (len(x), 240) and not xp[:, 160:].any()


def test_model_checkpoint_matches_reference(golden_dir):
    for case in json.load(open(os.path.join(golden_dir, 'ref_checkpoint.json'))):
        lr, count, saved = case['lr0'], 0, 0
        for ep, exp in zip(range(1, len(case['series'])), case['trace']):
            save, lr, count = V.model_checkpoint({'Train loss': case['series'][:ep + 1]}, lr, count)
            saved += save
            assert [saved, lr, -1 if count is None else count] == exp
            if count is None: break


def test_layout_and_init():
    lay, n = V.param_layout([80, 40, 20, 10], 400)
    assert n == 73300
    assert V.n_params([80, 40, 10], 80) == 20740        # cfg-1
    assert V.n_params([1024, 512, 256, 64], 400) == 2183184
    assert [x[0] for x in lay[:2]] == ['enc0.kernel', 'enc0.bias'] and lay[-1][0] == 'out.bias'
    p = V.init_params([80, 40, 10], 80, 0)
    assert p.dtype == np.float32 and np.array_equal(p, V.init_params([80, 40, 10], 80, 0))
    Wm, bm = V.unflatten(p, [80, 40, 10], 80)[2]
    assert Wm.shape == (40, 10) and not bm.any()


@pytest.mark.parametrize('oe', V.OE_TYPES)
def test_numpy_backward_vs_torch_autograd(oe):
    """Two independent derivations of d(sum_i total_i)/d(theta) agree to 1e-9 in float64."""
    import torch
    from oracle import torch_twin as T
    rng = np.random.default_rng(3)
    FC, D, B = [24, 12, 6], 20, 33
    flat = V.init_params(FC, D, 1).astype(np.float64)
    X = rng.normal(0, 1, (B, D)); Xo = rng.normal(0.2, 1.3, (B, D)); X[:, 15:] = 0
    w = rng.uniform(.5, 2, B); wo = rng.uniform(.5, 2, B); eps = rng.normal(size=(B, 6))
    kw = dict(OE_type=oe, beta=0.7, lamb=1.9, margin=0.05)
    (r, k, o, t), g = V.losses_and_grad(flat, FC, D, X, w, Xo, wo, eps, **kw)
    vae = T.TorchVAE(flat, FC, D, torch.float64)
    tt = lambda a: torch.tensor(a, dtype=torch.float64)
    parts = T.get_losses(vae, tt(X), tt(w), tt(Xo), tt(wo), tt(eps), oe, 0.7, 1.9, 0.05)
    gt = torch.autograd.grad(parts[3].sum(), vae.vars, allow_unused=True)
    gt = np.concatenate([(torch.zeros_like(p) if q is None else q).numpy().ravel() for q, p in zip(gt, vae.vars)])
    for a, b in zip((r, k, o, t), parts):
        np.testing.assert_allclose(a, b.detach().numpy(), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(g, gt, rtol=1e-9, atol=1e-11)
    assert np.abs(g).max() > 1e-3 and (o > 0).any()


def test_clip_paths_and_adam():
    t = np.array([np.nan, np.inf, -np.inf, 2e6, -2e6, 1e6, 3.0])
    np.testing.assert_array_equal(V.clipv(t), [0, 0, 0, 1e6, -1e6, 1e6, 3.0])
    np.testing.assert_array_equal(V.clip_mask(t), [0, 0, 0, 0, 0, 1, 1])
    # lv = +-200: exp overflows the clip in fp32; losses stay finite and the gradient is masked (models.py:84)
    mu = np.zeros((2, 4), np.float32); lv = np.array([[200.] * 4, [-200.] * 4], np.float32)
    k = V.kld_loss(mu, lv)
    # fp32: exp(200) = inf -> clip_values maps NON-FINITE to 0 (not to 1e6); float64: 7e86 is finite -> clamps to 1e6
    assert np.isfinite(k).all() and k[0] == np.float32(-(1 + 200 - 0) / 2) and k[1] == np.float32(-(1 - 200) / 2)
    assert V.kld_loss(mu.astype(np.float64), lv.astype(np.float64))[0] == -(1 + 200 - 1e6) / 2
    # Keras Adam: epsilon outside the bias correction (F10); first step moves by ~lr*sign(g)
    g = np.array([1e-3, -5.0, np.nan])
    th, m, v = V.adam_step(np.zeros(3), np.zeros(3), np.zeros(3), g, 1, 1e-3)
    gs = np.array([1e-3, -5.0, 0.0])                                   # NaN gradient sanitised to 0 (models.py:185)
    alpha = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
    np.testing.assert_allclose(th, -alpha * 0.1 * gs / (np.sqrt(0.001 * gs * gs) + 1e-7), rtol=1e-12)
    assert th[2] == 0 and abs(th[1] - 1e-3) < 1e-8 and abs(th[0] + 1e-3) < 5e-6


def test_torch_twin_train_step_matches_numpy_trainer():
    import torch
    from oracle import torch_twin as T
    rng = np.random.default_rng(5)
    FC, D, B = [16, 8, 4], 12, 20
    flat = V.init_params(FC, D, 2)
    X = rng.normal(size=(B, D)).astype(np.float32); Xo = rng.normal(size=(B, D)).astype(np.float32)
    w = np.ones(B, np.float32); eps = rng.normal(size=(B, 4)).astype(np.float32)
    tr = V.Trainer(flat, FC, D, 'KLD', np.float64)
    vae = T.TorchVAE(flat, FC, D, torch.float64)
    tt = lambda a: torch.tensor(a, dtype=torch.float64)
    for _ in range(5):
        tr.step(X, w, Xo, w, eps, lr=1e-2, beta=1.0, lamb=1.0, margin=1.0)
        T.train_step(vae, tt(X), tt(w), tt(Xo), tt(w), tt(eps), 'KLD', 1.0, 1.0, 1.0, 1e-2)
    np.testing.assert_allclose(tr.theta, vae.flat(), rtol=1e-9, atol=1e-12)


def test_oracle_regression_vectors(golden_dir):
    g = np.load(os.path.join(golden_dir, 'oracle_train.npz'))
    FC, D = [int(v) for v in g['FC']], int(g['D'])
    for oe in V.OE_TYPES:
        out, gr = V.losses_and_grad(g['flat0'], FC, D, g['X'], g['w'], g['Xo'], g['wo'], g['eps'], OE_type=oe,
                                    beta=0.5, lamb=2.0, margin=0.3)
        np.testing.assert_allclose(np.stack(out), g['loss_' + oe], rtol=1e-12)
        np.testing.assert_allclose(gr, g['grad_' + oe], rtol=1e-10, atol=1e-14)


# ---- non-affine scalers: the restatement the CUDA kernels follow, pinned against scikit-learn / SciPy / NumPy themselves
def _jet_like(rng, N, D):
    X = (rng.gamma(2.0, 20.0, (N, D)) * (rng.uniform(size=(N, D)) > 0.35)).astype(np.float32)
    X[:, 1::4] *= np.where(rng.uniform(size=(N, D // 4)) < .5, -1, 1).astype(np.float32)
    return X


def test_ndtri_and_interp_restatements_match_scipy_numpy():
    import scipy.special as sp
    from oracle import scalers as Sc
    rng = np.random.default_rng(0)
    y = np.concatenate([rng.uniform(0, 1, 100000), 10.0 ** rng.uniform(-300, -1, 10000), 1 - 10.0 ** rng.uniform(-16, -1, 10000), [0, 1, 0.5, Sc.EXPM2, 1 - Sc.EXPM2]])
    a, b = Sc.ndtri(y), sp.ndtri(y)
    fin = np.isfinite(b)
    assert np.array_equal(a[~fin], b[~fin]) and np.mean(a == b) > 0.999
    np.testing.assert_allclose(a[fin], b[fin], rtol=2e-15, atol=0)
    xp = np.sort(rng.normal(size=1000)); xp[100:110] = xp[100]; fp = np.linspace(0, 1, 1000)
    x = np.concatenate([rng.normal(size=20000), xp[::7], [xp[0], xp[-1], -9, 9, np.nan]])
    np.testing.assert_array_equal(Sc.interp(x, xp, fp), np.interp(x, xp, fp))
    np.testing.assert_array_equal(Sc.interp(-x, -xp[::-1], -fp[::-1]), np.interp(-x, -xp[::-1], -fp[::-1]))


def test_quantile_and_power_restatements_match_sklearn():
    import warnings
    from sklearn import preprocessing
    from oracle import scalers as Sc
    rng = np.random.default_rng(1)
    X = _jet_like(rng, 12000, 16)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        qt = preprocessing.QuantileTransformer(output_distribution='normal', n_quantiles=10000, random_state=0).fit(X)   # utils.py:394
        Xn = np.concatenate([X[:3000], (X[:2000] * 1.7 + 3).astype(np.float32), np.full((1, 16), np.nan, np.float32)])
        ref = qt.transform(Xn.copy())
    assert ref.dtype == np.float32
    np.testing.assert_array_equal(Sc.quantile_normal_transform(Xn, qt.quantiles_, qt.references_), ref)              # bit-exact, NaN included
    pt = preprocessing.PowerTransformer().fit(X[:4000])                                                             # utils.py:396
    assert pt.lambdas_.dtype == np.float32
    np.testing.assert_array_equal(Sc.power_transform(X, pt.lambdas_, pt._scaler.mean_, pt._scaler.scale_), pt.transform(X.copy()))
