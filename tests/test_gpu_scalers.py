"""Non-affine scaler transforms on the device against scikit-learn's own transform of the same float32 array:
QuantileTransformer(normal) bit for bit (the one difference allowed is the last bit of the float64 logarithm in the 1e-7 tails),
PowerTransformer to float32 ulps (its log1p / expm1 run in float32 in NumPy and on the device alike, with different roundings)."""
import warnings
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu


def _jet_like(rng, N, D):
    X = (rng.gamma(2.0, 20.0, (N, D)) * (rng.uniform(size=(N, D)) > 0.35)).astype(np.float32)
    X[:, 1::4] *= np.where(rng.uniform(size=(N, D // 4)) < .5, -1, 1).astype(np.float32)
    return X


@pytest.mark.parametrize('D', [16, 400])
def test_quantile_transformer_bit_exact(D):
    from sklearn import preprocessing
    from atlas_vae_b200 import utils as U
    from conftest import record_parity
    rng = np.random.default_rng(1)
    X = _jet_like(rng, 20000, D)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        qt = preprocessing.QuantileTransformer(output_distribution='normal', n_quantiles=10000, random_state=0).fit(X)   # utils.py:394
        Xn = np.concatenate([X[:6000], (X[:3000] * 1.7 + 3).astype(np.float32), (X[:3000] * 0.01).astype(np.float32), np.full((1, D), np.nan, np.float32)])
        ref = qt.transform(Xn.copy())
    got = U.apply_scaler(Xn, 4, qt)
    assert got.dtype == np.float32 and got.shape == ref.shape
    same = (got == ref) | (np.isnan(got) & np.isnan(ref))
    ulp = np.abs(got.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64))[~same]
    record_parity('QuantileTransformer(normal) apply, D=%d: fraction of elements not bit-identical to sklearn' % D, 1 - same.mean(), 1e-6,
                  {'max_ulp_of_the_rest': int(ulp.max()) if ulp.size else 0})
    assert 1 - same.mean() <= 1e-6 and (ulp.size == 0 or ulp.max() <= 1)


def test_power_transformer_close():
    from sklearn import preprocessing
    from atlas_vae_b200 import utils as U
    from conftest import record_parity
    rng = np.random.default_rng(2)
    X = _jet_like(rng, 20000, 40)
    pt = preprocessing.PowerTransformer().fit(X[:5000])                                                              # utils.py:396
    ref = pt.transform(X.copy())
    got = U.apply_scaler(X, 4, pt)
    err = np.abs(got.astype(np.float64) - ref) / np.maximum(np.abs(ref), 1.0)
    record_parity('PowerTransformer(yeo-johnson) apply: max |err| / max(|ref|, 1)', err.max(), 5e-6, {'bit_identical_fraction': float(np.mean(got == ref))})
    assert err.max() <= 5e-6


def test_preprocess_with_quantile_scaler_matches_host_chain():
    """K0 (sort / slice) + QuantileTransformer on the device == jets_sorting + slice + sklearn transform (utils.py:89-119,411)."""
    from sklearn import preprocessing
    from atlas_vae_b200 import utils as U, synthetic as S
    raw, _ = S.make_jets('QCD', 5000, 100, seed=3, device='cuda')
    x0, _, _ = U.preprocess_device(raw, 20, 4, None, True)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        qt = preprocessing.QuantileTransformer(output_distribution='normal', n_quantiles=10000, random_state=0).fit(x0.cpu().numpy())
        ref = qt.transform(x0.cpu().numpy().copy())
    x1, _, _ = U.preprocess_device(raw, 20, 4, qt, True)
    got = x1.cpu().numpy()
    assert np.mean(got != ref) <= 1e-6
