"""train_model (OE-VAE/models.py:135-219) on the native batch loop: what the drop-in API does around the step kernels.
  * the native host-load loop == stepping the same batches from Python (bit-identical parameters)
  * a second train_model call starts from a FRESH Adam (models.py:138), as the reference's does
  * float16 host rows + scaler on the device == the host-scaled float32 sample (bit-identical), and oevae_ingest_f16 == sklearn
  * lamb == 0: the OoD arrays are not needed (same parameters with garbage in them)
  * resident epochs (generator OFF) == re-uploading every epoch; metrics_every does not change the parameters"""
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

from oracle import vae_numpy as V, philox


def _sample(n, D, seed, dtype=np.float32):
    rng = np.random.default_rng(seed)
    X = rng.normal(0, 0.4, (n, D)).astype(np.float32)
    X[:, D // 2:] *= (rng.uniform(size=(n, D - D // 2)) > 0.4)
    return {'constituents': X.astype(dtype), 'weights': rng.uniform(0.5, 1.5, n).astype(np.float32)}


def _model(FC, D, seed=0, **kw):
    from atlas_vae_b200.models import VariationalAutoEncoder
    return VariationalAutoEncoder(FC, D, seed=seed, precision='bf16x3', **kw)


FC, D = [80, 40, 20, 10], 400


def test_native_loop_equals_python_stepping_and_oracle():
    from atlas_vae_b200.models import train_model
    n, bs = 2300, 500                                             # short last batch (300 rows)
    bkg, val = _sample(n, D, 1), _sample(700, D, 2)
    m = _model(FC, D)
    th0 = m.dev.get_params().copy()
    hist = train_model(m, [(bkg, bkg)], [(val, val)], 'KLD', 3, bs, 1.0, 0, 0, 1e-3, verbose=False)
    th_native = m.dev.get_params()
    # the same batches stepped from Python on a second handle
    m2 = _model(FC, D); m2.dev.set_params(th0)
    X = torch.as_tensor(bkg['constituents']).cuda(); w = torch.as_tensor(bkg['weights']).cuda()
    for ep in range(3):
        for lo in range(0, n, bs):
            m2.dev.train_step(X[lo:lo + bs], w[lo:lo + bs], None, None, None, 1e-3, 1.0, 0., 0.)
    assert np.array_equal(th_native, m2.dev.get_params())
    assert len(hist['Train loss']) == 3 and hist['Train loss'][2] < hist['Train loss'][0]
    # float64 oracle with the spec noise of the fixed mode (seed 0: the same eps rows for every batch)
    tr = V.Trainer(th0, FC, D, 'KLD')
    eps = philox.normal_eps(0, bs, FC[-1])
    tot = 0.0
    for ep in range(3):
        tot = 0.0
        for lo in range(0, n, bs):
            xb, wb = bkg['constituents'][lo:lo + bs], bkg['weights'][lo:lo + bs]
            o = tr.step(xb, wb, None, None, eps[:len(xb)], lr=1e-3, beta=1.0)
            tot += float(np.sum(o[3]))
    from test_gpu_tensorcore import _close
    _close([hist['Train loss'][-1]], [tot / n], 1e-3, 1e-6, 'train_model epoch loss vs oracle')


def test_second_call_starts_from_fresh_adam():
    from atlas_vae_b200.models import train_model
    bkg, val = _sample(1500, D, 3), _sample(500, D, 4)
    a = _model(FC, D); th0 = a.dev.get_params().copy()
    train_model(a, [(bkg, bkg)], [(val, val)], 'KLD', 2, 500, 1.0, 0, 0, 1e-3, verbose=False)
    th_a = a.dev.get_params()
    b = _model(FC, D)
    train_model(b, [(bkg, bkg)], [(val, val)], 'KLD', 1, 500, 1.0, 0, 0, 1e-3, verbose=False)      # leaves m, v, t behind
    _, _, t_b = b.dev.get_opt_state(); assert t_b == 3
    b.dev.set_params(th0)
    train_model(b, [(bkg, bkg)], [(val, val)], 'KLD', 2, 500, 1.0, 0, 0, 1e-3, verbose=False)
    assert np.array_equal(b.dev.get_params(), th_a)               # stale moments / step count would change every parameter
    _, _, t_b = b.dev.get_opt_state(); assert t_b == 6


def test_ingest_f16_matches_sklearn():
    from sklearn import preprocessing
    from atlas_vae_b200 import _ffi
    from atlas_vae_b200._ffi import ptr, check
    rng = np.random.default_rng(5)
    n = 3001
    x16 = (rng.normal(0, 30, (n, D)) * (rng.uniform(size=(n, D)) > 0.3)).astype(np.float16)
    x32 = x16.astype(np.float32)
    dev = torch.device('cuda', 0)
    t16 = torch.as_tensor(x16).to(dev)
    for name in ('MaxAbsScaler', 'RobustScaler', None):
        out = torch.empty((n, D), dtype=torch.float32, device=dev)
        c_t = s_t = None
        if name:
            sc = getattr(preprocessing, name)().fit(x32)
            ref = sc.transform(x32.copy())
            s_t = torch.as_tensor(np.asarray(sc.scale_, np.float64)).to(dev)
            if name == 'RobustScaler': c_t = torch.as_tensor(np.asarray(sc.center_, np.float64)).to(dev)
        else:
            ref = x32
        check(_ffi.lib().oevae_ingest_f16(ptr(t16), ptr(out), n, D, ptr(c_t), ptr(s_t), 0, torch.cuda.current_stream().cuda_stream))
        assert ref.dtype == np.float32
        np.testing.assert_array_equal(out.cpu().numpy(), ref)    # bit-exact with sklearn's float32 transform
    # ragged / unaligned: odd column count and an offset view
    x16b = x16[:77, :37].copy(); t = torch.as_tensor(x16b).to(dev); out = torch.empty((77, 37), dtype=torch.float32, device=dev)
    check(_ffi.lib().oevae_ingest_f16(ptr(t), ptr(out), 77, 37, None, None, 0, torch.cuda.current_stream().cuda_stream))
    np.testing.assert_array_equal(out.cpu().numpy(), x16b.astype(np.float32))


def test_fp16_host_sample_with_device_scaler_equals_host_scaling():
    from sklearn import preprocessing
    from atlas_vae_b200.models import train_model
    raw = _sample(2100, D, 6)
    raw['constituents'] = (raw['constituents'] * 50).astype(np.float16)          # what the files hold
    val = _sample(400, D, 7); val['constituents'] = (val['constituents'] * 50).astype(np.float16)
    scaler = preprocessing.MaxAbsScaler().fit(raw['constituents'].astype(np.float32))
    host = lambda s: {'constituents': scaler.transform(s['constituents'].astype(np.float32)), 'weights': s['weights']}
    a, b = _model(FC, D), _model(FC, D)
    st = {}
    ha = train_model(a, [(raw, raw)], [(val, val)], 'KLD', 2, 500, 1.0, 0, 0, 1e-3, verbose=False, scaler=scaler, stats=st)
    hb = train_model(b, [(host(raw), host(raw))], [(host(val), host(val))], 'KLD', 2, 500, 1.0, 0, 0, 1e-3, verbose=False)
    assert np.array_equal(a.dev.get_params(), b.dev.get_params())
    assert ha['Train loss'] == hb['Train loss'] and ha['Valid loss'] == hb['Valid loss']
    # the load was uploaded once (epoch 0), as float16: (2100 x 400 x 2 + weights) + the validation sample per epoch-0 upload
    assert st['h2d_bytes'] <= 2100 * (D * 2 + 4) + 400 * (D * 2 + 4) + 64 and st['steps'] == 10


def test_lamb_zero_ignores_ood_and_metrics_every_is_cosmetic():
    from atlas_vae_b200.models import train_model
    bkg, val = _sample(1700, D, 8), _sample(300, D, 9)
    junk = {'constituents': np.full((1700, D), np.nan, np.float32), 'weights': np.full(1700, np.nan, np.float32)}
    a, b, c = _model(FC, D), _model(FC, D), _model(FC, D)
    train_model(a, [(bkg, bkg)], [(val, val)], 'KLD', 2, 400, 1.0, 0, 0, 1e-3, verbose=False)
    train_model(b, [(bkg, junk)], [(val, val)], 'KLD', 2, 400, 1.0, 0, 0, 1e-3, verbose=False, metrics_every=1)
    assert np.array_equal(a.dev.get_params(), b.dev.get_params())
    # generator ON (two loads): every epoch re-uploads; same parameters as one load holding both halves
    h1 = {k: v[:800] for k, v in bkg.items()}; h2 = {k: v[800:1600] for k, v in bkg.items()}
    both = {k: v[:1600] for k, v in bkg.items()}
    d = _model(FC, D)
    train_model(c, [(h1, h1), (h2, h2)], [(val, val)], 'KLD', 2, 400, 1.0, 0, 0, 1e-3, verbose=False, metrics_every=3)
    train_model(d, [(both, both)], [(val, val)], 'KLD', 2, 400, 1.0, 0, 0, 1e-3, verbose=False)
    assert np.array_equal(c.dev.get_params(), d.dev.get_params())


def test_oe_training_through_train_model_matches_oracle():
    """cfg-3 through the drop-in API: paired OoD sample, OE_type KLD, lamb != 0 -> history keys and epoch means vs the oracle."""
    from atlas_vae_b200.models import train_model
    n, bs = 1200, 400
    bkg, ood, val = _sample(n, D, 10), _sample(n, D, 11), _sample(300, D, 12)
    m = _model(FC, D); th0 = m.dev.get_params().copy()
    hist = train_model(m, [(bkg, ood)], [(val, val)], 'KLD', 2, bs, 0.8, 1.5, 0.3, 1e-3, verbose=False)
    assert set(hist) == {'MSE', 'KLD', 'OE', 'Train loss', 'Valid loss'}
    tr = V.Trainer(th0, FC, D, 'KLD')
    eps = philox.normal_eps(0, bs, FC[-1])
    for ep in range(2):
        acc = np.zeros(4)
        for lo in range(0, n, bs):
            o = tr.step(bkg['constituents'][lo:lo + bs], bkg['weights'][lo:lo + bs], ood['constituents'][lo:lo + bs], ood['weights'][lo:lo + bs],
                        eps, lr=1e-3, beta=0.8, lamb=1.5, margin=0.3)
            acc += [float(np.sum(q)) for q in o]
    from test_gpu_tensorcore import _close
    got = [hist['MSE'][-1], hist['KLD'][-1], hist['OE'][-1], hist['Train loss'][-1]]
    _close(got, acc / n, 1e-3, 1e-5, 'train_model OE epoch means vs oracle')
