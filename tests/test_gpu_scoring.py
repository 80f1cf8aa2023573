"""GPU parity tests of the scoring pass, preprocessing (K0) and cut + histogram (K5) against the oracle and the
reference-generated golden fixtures.  Integer / index / byte results are bit-exact; float metrics carry the
tolerance written next to each check."""
import os
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

from oracle import vae_numpy as V, scoring as S, preprocess as PP


from test_gpu_tensorcore import _close          # records the achieved error in gpurun_out/r02_parity.jsonl


def test_pair_metrics_match_reference(golden_dir):
    """loss_function (utils.py:346-375) on the reference's own golden P/Q incl. 0/0, x/0 and sign flips.
    fp32 sums in a different order than NumPy's pairwise sum: rel 2e-5."""
    from atlas_vae_b200 import utils as U
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    for m in ['MSE', 'MAE', 'MARE', 'KLD', 'JSD', 'X-S']:
        got = U.loss_function(g['P'], g['Q'], 4, m, multiloss=False)
        ref = g['m_' + m]
        assert got.dtype == np.float32 and got.shape == ref.shape
        fin = np.isfinite(ref)
        assert np.array_equal(np.isfinite(got), fin), m
        scale = np.abs(ref[fin]).max()
        big = np.abs(ref) > 1e30                                # rows dominated by nan_to_num(+-inf) = +-FLT_MAX
        _close(got[fin & ~big], ref[fin & ~big], 2e-5, 2e-6 * max(scale, 1) if scale < 1e30 else 1e-3, m)
        assert np.array_equal(np.sign(got[big]), np.sign(ref[big]))
    kept = U.filtering(np.arange(len(g['P'])), g['P'].copy(), g['filt_bad'].copy(), {'m': np.arange(len(g['P']))})[0]
    np.testing.assert_array_equal(kept, g['filt_kept'])
    from atlas_vae_b200 import cuts as Cc
    for i in range(5):
        np.testing.assert_allclose(Cc.loss_mapping(g['map_in%d' % i]), g['map_out%d' % i], rtol=2e-7, atol=0)


@pytest.mark.parametrize('n_iter', [1, 3])
def test_score_vs_oracle(n_iter):
    from atlas_vae_b200.engine import DeviceVAE
    rng = np.random.default_rng(2)
    FC, D, N = [80, 40, 20, 10], 400, 1500
    flat = V.init_params(FC, D, 4)
    X = rng.normal(0, 0.4, (N, D)).astype(np.float32); X[:, 300:] = 0
    eps = rng.normal(size=(n_iter, N, 10)).astype(np.float32)
    dev = DeviceVAE(FC, D, 'KLD', 'fp32', 'external', 0); dev.set_params(flat)
    out = dev.score(X, n_iter, eps, metrics=('MSE', 'MAE', 'KLD', 'JSD', 'Latent'), want_xhat=True)
    xhat = S.predict(flat, FC, D, X, list(eps), np.float64)
    _close(out['xhat'].cpu().numpy(), xhat, 1e-3, 2e-5, 'xhat')
    ref = S.score_metrics(X.astype(np.float64), xhat.astype(np.float64), ('MSE', 'MAE'))
    _close(out['MSE'].cpu().numpy(), ref['MSE'], 1e-3, 1e-6, 'MSE')
    _close(out['MAE'].cpu().numpy(), ref['MAE'], 1e-3, 1e-6, 'MAE')
    # KLD / JSD contain log2(p/q) of near-cancelling terms: compare with the oracle evaluated on the DEVICE's
    # own xhat in float32 (checks the metric arithmetic, not the network twice)
    ref2 = S.score_metrics(X, out['xhat'].cpu().numpy(), ('KLD', 'JSD'))
    for m in ('KLD', 'JSD'):
        scale = np.abs(ref2[m]).max()
        _close(out[m].cpu().numpy(), ref2[m], 1e-3, 1e-4 * scale, m)
    _close(out['Latent'].cpu().numpy(), S.latent_score(flat, FC, D, X, np.float64), 1e-3, 1e-6, 'Latent')
    assert out['finite'].all()


def test_model_surface_and_score_sample(tmp_path):
    """VariationalAutoEncoder call / encoder / losses / predict / save+load, get_losses, score_sample."""
    from atlas_vae_b200 import VariationalAutoEncoder, get_losses
    from atlas_vae_b200 import cuts as Cc, utils as U
    rng = np.random.default_rng(3)
    FC, D, N = [32, 16, 8], 40, 600
    vae = VariationalAutoEncoder(FC, D, seed=0)
    flat = vae.dev.get_params()
    assert np.array_equal(flat, V.init_params(FC, D, 0))            # same initialiser spec as the oracle's
    X = rng.normal(0, 0.5, (N, D)).astype(np.float32)
    from oracle import philox
    eps = philox.normal_eps(0, N, 8)
    rec = vae(X)
    _close(rec, S.predict(flat, FC, D, X, [eps], np.float64), 1e-3, 2e-5, 'vae(X)')
    _close(vae.losses[0], S.latent_score(flat, FC, D, X, np.float64), 1e-3, 1e-6, 'vae.losses')
    mu, lv, z = vae.encoder(X)
    assert mu.shape == (N, 8) and np.array_equal(vae.predict(X, batch_size=100), rec)
    tw = vae.trainable_weights
    assert len(tw) == 2 * (2 * 2 + 3) and tw[0].shape == (D, 32) and tw[-1].shape == (D,) and hasattr(tw[0], 'numpy')
    p = str(tmp_path / 'model.h5'); vae.save_weights(p)
    vae2 = VariationalAutoEncoder(FC, D, seed=0, init_seed=9); vae2.load_weights(p)
    assert np.array_equal(vae2.dev.get_params(), flat)
    data = ({'constituents': X, 'weights': np.ones(N, np.float32)}, {'constituents': X[::-1].copy(), 'weights': np.ones(N, np.float32)})
    parts = get_losses(vae, data, 'KLD', 1.0, 2.0, 0.5)
    ref, _ = V.losses_and_grad(flat, FC, D, X, np.ones(N), X[::-1], np.ones(N), eps, OE_type='KLD', beta=1.0, lamb=2.0, margin=0.5, want_grad=False)
    for a, b in zip(parts, ref): _close(a.numpy(), b, 1e-3, 1e-6, 'get_losses')
    scores, fin = Cc.score_sample(vae, X, metrics=('Latent', 'MAE'), normal_losses='ON')
    mae = S.score_metrics(X.astype(np.float64), S.predict(flat, FC, D, X, [eps], np.float64).astype(np.float64), ('MAE',))['MAE']
    _close(scores['MAE'], S.loss_mapping(mae), 1e-3, 1e-6, 'mapped MAE')
    assert fin.all() and np.array_equal(U.latent_loss(X, vae), np.asarray(vae.losses[0]))


def test_preprocess_bit_exact(golden_dir):
    """K0 vs the reference's jets_sorting / get_4v outputs: bit-exact (sort, gather, fp32 sums in NumPy's order)."""
    from atlas_vae_b200 import utils as U
    g = np.load(os.path.join(golden_dir, 'ref_preprocess.npz'))
    np.testing.assert_array_equal(U.jets_sorting(g['raw']), g['sorted'])
    np.testing.assert_array_equal(U.jets_sorting(g['raw16']), g['sorted16'])
    s = U.preprocess(g['raw'], 20, 4)
    np.testing.assert_array_equal(s['constituents'], g['sorted'][:, :80])
    np.testing.assert_array_equal(s['pt'], g['pt20']); np.testing.assert_array_equal(s['m'], g['m20'])
    fv = U.get_4v(g['sorted'][:, :80])
    np.testing.assert_array_equal(fv['pt_calo'], g['pt20']); np.testing.assert_array_equal(fv['m_calo'], g['m20'])
    x3 = U.preprocess(g['raw'], 20, 3)['constituents']
    np.testing.assert_array_equal(x3, PP.preprocess(g['raw'], 20, 3)[0])
    xp = U.preprocess(g['raw'], 60, 4)['constituents']                       # zero padding beyond n_raw
    assert xp.shape == (len(g['raw']), 240) and not xp[:, 160:].any()
    np.testing.assert_array_equal(xp[:, :160], g['sorted'])
    assert U.preprocess(g['raw'][:0], 20, 4)['constituents'].shape == (0, 80)    # empty input


def test_preprocess_scalers_and_large():
    """Affine scalers are bit-exact with sklearn's own transform; 200-constituent jets from the generator."""
    from sklearn import preprocessing
    from atlas_vae_b200 import utils as U, synthetic as Sy
    raw, _ = Sy.make_jets('top', 3000, 200, seed=5)
    raw = raw.numpy()
    x_ref, pt_ref, m_ref = PP.preprocess(raw, 100, 4)
    s = U.preprocess(raw, 100, 4)
    np.testing.assert_array_equal(s['constituents'], x_ref)
    np.testing.assert_array_equal(s['pt'], pt_ref); np.testing.assert_array_equal(s['m'], m_ref)
    for scaler in (preprocessing.MaxAbsScaler(), preprocessing.RobustScaler()):
        scaler.fit(x_ref[:1000])
        with np.errstate(all='ignore'):
            ref = scaler.transform(x_ref)
        got = U.preprocess(raw, 100, 4, scaler)['constituents']
        ok = np.isfinite(ref)
        np.testing.assert_array_equal(got[ok], ref[ok])
        np.testing.assert_array_equal(U.apply_scaler(x_ref, 4, scaler)[ok], ref[ok])


@pytest.mark.parametrize('n_raw', [200, 50, 32, 7])
def test_preprocess_partial_sort_cases(n_raw):
    """K0 compacts the slots with pT != 0, appends the pT == 0 slots in descending index order and sorts only the compacted part.
    Cases that exercise it against the oracle's argsort(pT)[::-1] (utils.py:564-571), bit-exact: every slot filled, no slot filled,
    one slot, exact pT ties, pT == 0 slots that still carry E / pz and sit between filled slots, -0.0 momenta, n_raw that is not a
    multiple of the warp size."""
    from atlas_vae_b200 import utils as U
    rng = np.random.default_rng(41 + n_raw)
    N = 64
    raw = rng.normal(0, 5, (N, n_raw, 4)).astype(np.float32)
    raw[1] = 0                                                            # empty jet
    raw[2, 1:] = 0                                                        # one constituent
    raw[3, :, 1:3] = raw[3, 0, 1:3]                                       # every pT equal: order by descending index
    raw[4, ::3, 1:3] = 0                                                  # pT == 0 with E, pz != 0 between filled slots
    raw[5, ::2, 1] = -0.0; raw[5, ::2, 2] = 0.0
    for j in range(6, N):                                                 # zero padding of random length, some jets full
        k = int(rng.integers(0, n_raw + 1)) if j % 4 else n_raw
        raw[j, k:] = 0
        if k > 2: raw[j, 1, 1:3] = raw[j, 0, 1:3]                         # one tie per jet
    raw = raw.reshape(N, 4 * n_raw)

    def sort_ref(r):                                                      # the device's documented tie rule: reverse of a STABLE argsort
        j = np.float32(r).reshape(len(r), -1, 4)                          # (NumPy's default argsort is not stable; without ties both agree)
        pt = np.sqrt(j[:, :, 1] ** 2 + j[:, :, 2] ** 2)
        order = np.argsort(pt, axis=-1, kind='stable')[:, ::-1]
        return np.take_along_axis(j, order[:, :, None], axis=1).reshape(len(r), -1)

    for n_const in sorted({n_raw, max(1, n_raw // 2), n_raw + 3}):
        x4 = PP.select_constituents(sort_ref(raw), n_const)
        pt_ref, m_ref = PP.jet_pt_mass(x4)
        s = U.preprocess(raw, n_const, 4)
        np.testing.assert_array_equal(s['constituents'], x4)
        np.testing.assert_array_equal(s['pt'], pt_ref); np.testing.assert_array_equal(s['m'], m_ref)
    np.testing.assert_array_equal(U.jets_sorting(raw.astype(np.float16)), sort_ref(raw.astype(np.float16)))


def test_hist_and_cuts_bit_exact(golden_dir):
    """K5 vs np.histogram on the reference's get_idx edges; cut masks vs the reference's make_cut selections."""
    from atlas_vae_b200 import cuts as Cc
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    for tag, bs in [('m', 2.5), ('pt', 10)]:
        var = g['var_' + tag]
        sample = {tag: var, 'weights': np.ones(len(var), np.float32), 'JZW': np.zeros(len(var), np.float32)}
        edges, c, ws = Cc.mass_histograms(sample, tag, bs)
        np.testing.assert_array_equal(edges, g['edges_' + tag])
        np.testing.assert_array_equal(c[1], g['hist_' + tag]); assert c[0].sum() == 0
        np.testing.assert_array_equal(ws[1], c[1].astype(np.float64))
    # mixed classes, weights, a cut, NaN scores / NaN masses / out-of-range values
    rng = np.random.default_rng(8)
    N = 200000
    m = rng.gamma(3.0, 30.0, N).astype(np.float32); m[5] = np.nan; m[6] = -1.0
    jzw = np.where(rng.uniform(size=N) < 0.2, -1, rng.integers(0, 9, N)).astype(np.float32)
    w = rng.uniform(0.1, 3, N).astype(np.float32)
    score = rng.beta(2, 5, N).astype(np.float32); score[::1000] = np.nan
    edges = S.bin_edges(m[np.isfinite(m)], 2.5)
    edges[3] = np.float64(m[100])                              # a value exactly on an inner edge
    edges = np.sort(edges)
    cut = np.float32(0.31)
    sample = {'m': m, 'weights': w, 'JZW': jzw}
    _, c, ws = Cc.mass_histograms(sample, 'm', 2.5, score, cut, edges=edges)
    with np.errstate(invalid='ignore'):
        cref, wref = S.class_histograms(np.nan_to_num(m, nan=-5.0), w, jzw, edges, S.cut_mask(score, cut))
    np.testing.assert_array_equal(c, cref)
    np.testing.assert_allclose(ws, wref, rtol=1e-12, atol=1e-9 * wref.sum())
    # reference make_cut selections (plots.py:250-259) through the product's host logic
    rates = Cc.get_rates(g['roc_y'], g['roc_score'], g['roc_w'])
    samp = {'weights': g['roc_w'], 'idx': np.arange(len(g['roc_y']))}
    np.testing.assert_array_equal(Cc.make_cut(g['roc_y'], g['roc_score'], samp, rates, 'MAE', 'gain')['idx'], g['kept_gain'])
    for eff in [1e-1, 1e0, 1e1]:
        np.testing.assert_array_equal(Cc.make_cut(g['roc_y'], g['roc_score'], samp, rates, 'MAE', 'bkg_eff', eff)['idx'], g['kept_eff_%g' % eff])
    # the device cut mask agrees with the host one: histogram count after the cut == number of kept jets in range
    s2 = {'m': np.abs(g['roc_score']) * 100, 'weights': g['roc_w'], 'JZW': np.where(g['roc_y'] == 1, 0, -1).astype(np.float32)}
    thr = Cc.cut_value(g['roc_y'], rates, g['roc_w'], 'gain')
    e2, c2, _ = Cc.mass_histograms(s2, 'm', 2.5, g['roc_score'], thr)
    assert c2.sum() == len(g['kept_gain'])
