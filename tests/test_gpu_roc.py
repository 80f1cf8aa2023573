"""Device threshold search: get_rates / best_threshold / make_cut / bump_scan cut grid (plots.py:228-259,272-283)
against scikit-learn's own roc_curve -- BIT-EXACT float64 fpr / tpr / thresholds -- and the reference-generated golden vectors; the
normalize / density weight scalings of the cut histograms (plots.py:706-715) against np.histogram."""
import os
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

from oracle import scoring as S


def _roc_sklearn(y, s, w):
    from sklearn import metrics
    return metrics.roc_curve(y, s, pos_label=0, sample_weight=w)


def _case(n, seed, ties, weights):
    rng = np.random.default_rng(seed)
    y = (rng.uniform(size=n) < 0.8).astype(np.int64)                       # 1 = QCD (negative class), 0 = signal (pos_label)
    s = np.where(y == 0, rng.beta(3, 3, n), rng.beta(2, 6, n)).astype(np.float32)
    if ties: s = np.round(s, ties).astype(np.float32)                     # heavy ties
    if weights == 'none': w = None
    elif weights == 'ones': w = np.ones(n, np.float32)
    elif weights == 'int': w = rng.integers(1, 7, n).astype(np.float32)
    elif weights == 'float': w = rng.uniform(0.1, 3.0, n).astype(np.float32)
    else:                                                                  # float weights with zeros (dropped by sklearn before sorting)
        w = rng.uniform(0.1, 3.0, n).astype(np.float32); w[rng.uniform(size=n) < 0.05] = 0
    return y, s, w


@pytest.mark.parametrize('n,ties,weights', [(1, 0, 'ones'), (2, 0, 'float'), (7, 1, 'float'), (4097, 2, 'int'), (100_000, 3, 'float'), (100_000, 0, 'zeros'),
                                            (1_000_000, 3, 'ones'), (1_000_000, 4, 'float'), (1_000_000, 0, 'none'), (300_001, 2, 'zeros')])
def test_roc_bit_exact_vs_sklearn(n, ties, weights):
    from atlas_vae_b200 import cuts as Cc
    y, s, w = _case(n, 1000 + n % 97, ties, weights)
    if n <= 2: y[:] = [0, 1][:n]
    roc = Cc.roc_curve_device(y, s, w)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        fpr, tpr, thr = _roc_sklearn(y, s, w)
    got = [t.cpu().numpy() for t in (roc.fpr, roc.tpr, roc.thresholds)]
    assert got[0].dtype == np.float64 and len(got[0]) == len(fpr), (len(got[0]), len(fpr))
    np.testing.assert_array_equal(got[2], thr)
    np.testing.assert_array_equal(got[0], fpr)
    np.testing.assert_array_equal(got[1], tpr)


def test_roc_parallel_and_sequential_cumsum_agree(monkeypatch):
    """Integer weights: the parallel scan (default for integer weights) and the sequential np.cumsum-order chain give the same bits."""
    from atlas_vae_b200 import cuts as Cc
    y, s, w = _case(500_000, 5, 3, 'int')
    out = {}
    for mode in ('par', 'seq'):
        monkeypatch.setenv('OEVAE_ROC_CUMSUM', mode)
        r = Cc.roc_curve_device(y, s, w)
        out[mode] = [t.cpu().numpy() for t in (r.fpr, r.tpr, r.thresholds)]
    for a, b in zip(out['par'], out['seq']): np.testing.assert_array_equal(a, b)


def test_thresholds_and_selections_vs_reference_golden(golden_dir):
    """get_rates -> best_threshold / make_cut on the golden ROC sample: rates, thresholds and kept-index sets equal the outputs of the
    reference's own function bodies (tests/golden/make_golden.py)."""
    from atlas_vae_b200 import cuts as Cc
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    y, s, w = g['roc_y'], g['roc_score'], g['roc_w']
    fpr, tpr, thr = Cc.get_rates(y, s, w)
    np.testing.assert_array_equal(fpr, g['roc_fpr']); np.testing.assert_array_equal(tpr, g['roc_tpr']); np.testing.assert_array_equal(thr, g['roc_thr'])
    roc = Cc.roc_curve_device(y, s, w)
    effs = [1e-1, 1e0, 1e1]
    cuts = roc.cuts(effs)
    assert cuts['gain'][0] == float(g['thr_gain']) and cuts['gain'][1] == float(g['val_gain'])
    idx = np.arange(len(y))
    np.testing.assert_array_equal(idx[s > cuts['gain'][0]], g['kept_gain'])
    for e, c in zip(effs, cuts['bkg_eff']):
        assert c == Cc.cut_value(y, (fpr, tpr, thr), w, 'bkg_eff', e)
        np.testing.assert_array_equal(idx[s > c], g['kept_eff_%g' % e])


def test_device_cuts_vs_host_logic_large():
    """1M scores with ties: the device arg-reductions against the reference's NumPy expressions on the same curve; bump_scan's cut grid."""
    from atlas_vae_b200 import cuts as Cc
    y, s, w = _case(1_000_000, 77, 4, 'float')
    roc = Cc.roc_curve_device(y, s, w)
    rates = roc.rates()
    effs = [1e-4, 1e-3, 1e-2, 1e-1, 1e0, 1e1]
    cuts = roc.cuts(effs)
    for e, c in zip(effs, cuts['bkg_eff']):
        assert c == Cc.cut_value(y, rates, w, 'bkg_eff', e)
    thr_h, val_h = Cc.best_threshold(y, rates, w, 'gain')
    assert cuts['gain'] == (float(thr_h), float(val_h))
    # bump_scan (plots.py:272-283)
    fpr, tpr, thr = rates
    logit = lambda x: np.log10(x) - np.log10(1 - x)
    inverse_logit = lambda x: 1 / (1 + 10 ** (-x))
    x_min = 10 ** np.ceil(np.log10(np.min(fpr)))
    eff_val = np.append(100 * inverse_logit(np.linspace(logit(x_min / 100), -logit(x_min / 100), 100)), 100)
    idx = np.minimum(np.searchsorted(fpr, eff_val, side='right'), len(fpr) - 1)
    t_d, e_d = Cc.bump_scan_cuts(roc, 100)
    np.testing.assert_array_equal(t_d, np.take(thr, idx)); np.testing.assert_array_equal(e_d, np.take(fpr, idx))


def test_density_histograms_vs_numpy():
    """plots.py:706-715: weights *= 100 / sum(weights of the uncut sample); weights /= bin width; hist(variable, bins, weights)."""
    from atlas_vae_b200 import cuts as Cc, utils as U
    rng = np.random.default_rng(3)
    N = 300_000
    m = rng.gamma(3.0, 30.0, N).astype(np.float32)
    jzw = np.where(rng.uniform(size=N) < 0.2, -1, rng.integers(0, 9, N)).astype(np.float32)
    for wkind in ('ones', 'float'):
        w = np.ones(N, np.float32) if wkind == 'ones' else rng.uniform(0.1, 3, N).astype(np.float32)
        sample = {'m': m, 'weights': w, 'JZW': jzw}
        edges, c, dens = Cc.mass_histograms(sample, 'm', 2.5, normalize=True, density=True)
        bins = U.get_idx(np.max(m), bin_size=2.5, min_val=max(0, np.min(m)), integer=False, tuples=False)
        np.testing.assert_array_equal(edges, bins)
        for row, cond in ((0, jzw == -1), (1, jzw >= 0)):
            variable = np.float32(m[cond]); weights = w[cond].copy()
            weights *= 100 / np.sum(w)                                      # plots.py:707
            indices = np.searchsorted(bins, variable, side='right')         # plots.py:710-711
            weights /= np.take(np.diff(bins), np.minimum(indices, len(bins) - 1) - 1)
            assert weights.dtype == np.float32
            ref_c = np.histogram(variable, bins)[0]
            ref_w = np.histogram(variable, bins, weights=weights.astype(np.float64))[0]   # float64 sum of the reference's float32 weights
            np.testing.assert_array_equal(c[row], ref_c)
            np.testing.assert_allclose(dens[row], ref_w, rtol=1e-12, atol=1e-12 * ref_w.sum())
            # what pylab.hist itself returns: np.histogram differences a float32 cumulative sum over 65536-element blocks, which
            # carries ~1e-3 of float32 rounding per bin; the device value is the float64 sum of the same float32 weights
            ref32 = np.histogram(variable, bins, weights=weights)[0]
            np.testing.assert_allclose(dens[row], ref32, rtol=5e-3, atol=1e-5 * ref32.max())


def test_generate_cuts_single_rank():
    from atlas_vae_b200 import cuts as Cc
    y, s, w = _case(200_000, 9, 0, 'ones')
    rng = np.random.default_rng(4)
    sample = {'m': rng.gamma(3.0, 30.0, len(y)).astype(np.float32), 'pt': rng.uniform(450, 1200, len(y)).astype(np.float32),
              'weights': w, 'JZW': np.where(y == 0, -1, 0).astype(np.float32)}
    res = Cc.generate_cuts(y, sample, s, 'MAE')
    rates = Cc.get_rates(y, s, w)
    assert set(res) == {'bkg_eff_%s' % format(e, '1.0e') for e in [1e-4, 1e-3, 1e-2, 1e-1, 1e0, 1e1]} | {'best_gain'}
    for e in [1e-2, 1e0, 1e1]:
        r = res['bkg_eff_' + format(e, '1.0e')]
        thr = Cc.cut_value(y, rates, w, 'bkg_eff', e)
        assert r['loss_cut'] == float(thr) and r['n_pass'] == int((s > thr).sum())
        e1, c1, d1 = r['m_cut']
        sel = s > thr
        ref_edges = Cc.bin_edges(sample['m'][sel], 2.5)
        np.testing.assert_array_equal(e1, ref_edges)
        np.testing.assert_array_equal(c1[1], np.histogram(sample['m'][sel & (y == 1)], ref_edges)[0])
        np.testing.assert_array_equal(c1[0], np.histogram(sample['m'][sel & (y == 0)], ref_edges)[0])
