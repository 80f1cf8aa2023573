"""GPU parity of the FUSED tcgen05 training kernel (precision bf16x3, no OoD sample) against the float64 oracle.
Same stated tolerances as tests/test_gpu_tensorcore.py for bf16x3."""
import os
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

from oracle import vae_numpy as V, philox
from test_gpu_tensorcore import _close, _close_grad


def _dev(FC, D, noise='external', seed=0, oe='KLD'):
    from atlas_vae_b200.engine import DeviceVAE
    return DeviceVAE(FC, D, oe, 'bf16x3', noise, seed)


def _data(FC, D, B, seed=11):
    rng = np.random.default_rng(seed)
    X = rng.normal(0, 0.7, (B, D)).astype(np.float32); X[:, D // 2:] *= (rng.uniform(size=(B, D - D // 2)) > 0.5)
    w = rng.uniform(0.5, 1.5, B).astype(np.float32)
    eps = rng.normal(size=(B, FC[-1])).astype(np.float32)
    return X, w, eps


@pytest.mark.parametrize('FC,D,B', [([80, 40, 10], 80, 256), ([80, 40, 20, 10], 400, 777), ([80, 40, 20, 10], 400, 5000),
                                    ([33, 7, 5], 23, 1), ([33, 7, 5], 23, 63), ([100, 16], 130, 65), ([64, 32, 16], 64, 64 * 151 + 3),
                                    ([6, 8], 20, 100)])       # first decoder layer narrower than the latent space: no replicated latent rows (heads_rep = 0)
@pytest.mark.parametrize('oe', ['KLD', 'MSE'])
def test_fused_losses_and_gradient(FC, D, B, oe):
    X, w, eps = _data(FC, D, B)
    flat = V.init_params(FC, D, 3)
    dev = _dev(FC, D, oe=oe); dev.set_params(flat)
    pj = torch.empty((4, B), device='cuda'); sums = torch.zeros(4, dtype=torch.float64, device='cuda')
    grad = dev.loss_and_grad(X, w, None, None, eps, 0.7, 0., 0., pj, sums)
    out, gref = V.losses_and_grad(flat, FC, D, X, w, None, None, eps, OE_type=oe, beta=0.7)
    ref = np.stack(out)
    _close(pj.cpu().numpy(), ref, 1e-3, 1e-6, 'per-jet losses')
    _close(sums.cpu().numpy(), ref.sum(1), 1e-4, 1e-6, 'sums')
    _close_grad(grad.cpu().numpy(), gref, 1e-3, 1e-4, 'grad', outliers=0.08 if B == 1 else 2e-3)
    pj2 = dev.eval_losses(X, w, None, None, eps, 0.7, 0., 0.)
    assert torch.equal(pj2, pj)                                   # eval mode == forward half of the train kernel


def test_fused_is_the_path_that_runs():
    from atlas_vae_b200 import _ffi
    FC, D, B = [80, 40, 20, 10], 400, 500
    X, w, eps = _data(FC, D, B)
    dev = _dev(FC, D); dev.set_params(V.init_params(FC, D, 3))
    _ffi.profile(True)
    dev.train_step(X, w, None, None, eps, 1e-3, 1.0)
    torch.cuda.synchronize()
    names = [r[0] for r in _ffi.profile_report()]
    _ffi.profile(False)
    assert 'fused_train' in names and 'adam' in names and not any(nm.startswith('gemm') for nm in names), names


def test_fused_adam_trajectory():
    FC, D, B = [80, 40, 10], 80, 300
    X, w, eps = _data(FC, D, B, 5)
    flat = V.init_params(FC, D, 0)
    dev = _dev(FC, D); dev.set_params(flat)
    tr = V.Trainer(flat, FC, D, 'KLD')
    Xd, wd, ed = [torch.as_tensor(v).cuda() for v in (X, w, eps)]
    curve, ref = [], []
    for s in range(20):
        sums = torch.zeros(4, dtype=torch.float64, device='cuda')
        dev.train_step(Xd, wd, None, None, ed, 1e-3, 0.5, 0., 0., None, sums)
        curve.append((sums / B).cpu().numpy())
        o = tr.step(X, w, None, None, eps, lr=1e-3, beta=0.5)
        ref.append([float(np.mean(q)) for q in o])
    _close(np.array(curve), np.array(ref), 1e-3, 1e-6, '20-step loss curve')
    th = dev.get_params()
    assert np.mean(np.abs(th - tr.theta) > 1e-3 * np.abs(tr.theta) + 5e-4) < 2e-3


def test_fused_philox_noise():
    """In-kernel Philox noise == the oracle's spec stream (fixed mode: same eps every call)."""
    FC, D, B = [80, 40, 10], 80, 200
    X, w, _ = _data(FC, D, B)
    flat = V.init_params(FC, D, 1)
    dev = _dev(FC, D, noise='fixed', seed=77); dev.set_params(flat)
    pj = dev.eval_losses(X, w, None, None, None, 1.0, 0., 0.)
    eps = philox.normal_eps(77, B, 10)
    out, _ = V.losses_and_grad(flat, FC, D, X, w, None, None, eps, OE_type='KLD', beta=1.0, want_grad=False)
    _close(pj.cpu().numpy(), np.stack(out), 1e-3, 1e-5, 'losses with Philox eps')
    assert torch.equal(dev.eval_losses(X, w, None, None, None, 1.0, 0., 0.), pj)


@pytest.mark.parametrize('N', [1, 63, 1500, 148 * 64 * 2 + 37])
def test_fused_score_vs_oracle(N):
    """Anomaly scoring (vae.py:163-167, utils.py:346-386) through the SCORE instantiation of the fused forward kernel:
    ragged N, several tiles per CTA.  X_hat / MSE / MAE / Latent vs the float64 oracle at rel 1e-3.  KLD / JSD hold
    log2(p/q) of near-cancelling ratios, so (as in test_gpu_scoring.test_score_vs_oracle) they are compared with the
    oracle's metric evaluated on the DEVICE's own X_hat: rel 1e-3 + 1e-4 of the vector's scale (SFU lg2 / reciprocal)."""
    from atlas_vae_b200.engine import DeviceVAE
    from oracle import scoring as S
    rng = np.random.default_rng(5)
    FC, D = [80, 40, 20, 10], 400
    flat = V.init_params(FC, D, 4)
    flat = (flat + rng.normal(0, 0.01, flat.shape)).astype(np.float32)     # init_params leaves the head / output biases at zero
    X = rng.normal(0, 0.4, (N, D)).astype(np.float32); X[:, 300:] = 0
    eps = rng.normal(size=(1, N, 10)).astype(np.float32)
    dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'external', 0); dev.set_params(flat)
    out = {k: v.cpu().numpy() for k, v in dev.score(X, 1, eps, metrics=('MSE', 'MAE', 'KLD', 'JSD', 'Latent'), want_xhat=True).items()}
    slim = {k: v.cpu().numpy() for k, v in dev.score(X, 1, eps, metrics=('MAE', 'Latent')).items()}     # no X_hat, two metrics
    np.testing.assert_array_equal(slim['MAE'], out['MAE']); np.testing.assert_array_equal(slim['Latent'], out['Latent'])
    n_ref = min(N, 2000)                                           # float64 oracle on a prefix (it is slow), device X_hat on all rows
    xhat = S.predict(flat, FC, D, X[:n_ref], [eps[0, :n_ref]], np.float64)
    _close(out['xhat'][:n_ref], xhat, 1e-3, 1e-4, 'xhat')           # abs floor: 2^-16-relative operand split over 8 layers of O(1) sums
    ref = S.score_metrics(X[:n_ref].astype(np.float64), xhat, ('MSE', 'MAE'))
    for m in ('MSE', 'MAE'):
        _close(out[m][:n_ref], ref[m], 1e-3, 1e-6, m)
    _close(out['Latent'][:n_ref], S.latent_score(flat, FC, D, X[:n_ref], np.float64), 1e-3, 1e-6, 'Latent')
    assert out['finite'].all()
    ref2 = S.score_metrics(X, out['xhat'], ('MSE', 'MAE', 'KLD', 'JSD'))
    for m in ('MSE', 'MAE'):
        _close(out[m], ref2[m], 1e-4, 1e-6, m + ' (device X_hat)')
    for m in ('KLD', 'JSD'):
        scale = max(np.abs(ref2[m]).max(), 1e-6)
        _close(out[m], ref2[m], 1e-3, 1e-4 * scale, m)


@pytest.mark.parametrize('FC,D,N', [([80, 40, 10], 80, 64 * 149 + 5), ([33, 7, 5], 23, 1000), ([100, 16], 130, 64 * 300 + 63), ([64, 32, 16], 64, 777), ([64, 32, 16], 64, 148 * 64 * 12 + 5),
                                    ([48, 24, 6], 12, 5000), ([80, 40, 20, 10], 1000, 700)])
def test_fused_score_shapes(FC, D, N):
    """Other layouts through the scoring schedule: fewer X chunks than chain stages (nX < 2n), one chunk, D <= 16 (one layer-0
    accumulator), 16 X chunks, decoder tiles that alias and tile regions too small for the two extra X ring slots; several tiles per
    CTA so that the tile queue wraps.  Same tolerances as test_fused_score_vs_oracle."""
    from atlas_vae_b200.engine import DeviceVAE
    from oracle import scoring as S
    rng = np.random.default_rng(31)
    L = FC[-1]
    flat = V.init_params(FC, D, 6)
    flat = (flat + rng.normal(0, 0.01, flat.shape)).astype(np.float32)     # init_params leaves the head / output biases at zero
    X = rng.normal(0, 0.4, (N, D)).astype(np.float32); X[:, (3 * D) // 4:] = 0
    eps = rng.normal(size=(1, N, L)).astype(np.float32)
    dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'external', 0); dev.set_params(flat)
    out = {k: v.cpu().numpy() for k, v in dev.score(X, 1, eps, metrics=('MSE', 'MAE', 'KLD', 'JSD', 'Latent'), want_xhat=True).items()}
    n_ref = min(N, 1000)
    xhat = S.predict(flat, FC, D, X[:n_ref], [eps[0, :n_ref]], np.float64)
    # abs floor: 2^-16-relative operand split per layer, so it scales with the magnitude the activations reach (narrow he-normal layers: ~10)
    _close(out['xhat'][:n_ref], xhat, 1e-3, 1e-4 * max(1.0, float(np.abs(xhat).max()) / 4), 'xhat')
    _close(out['Latent'][:n_ref], S.latent_score(flat, FC, D, X[:n_ref], np.float64), 1e-3, 1e-6, 'Latent')
    ref2 = S.score_metrics(X, out['xhat'], ('MSE', 'MAE', 'KLD', 'JSD'))
    for m in ('MSE', 'MAE'):
        _close(out[m], ref2[m], 1e-4, 1e-6, m + ' (device X_hat)')
    for m in ('KLD', 'JSD'):
        scale = max(np.abs(ref2[m]).max(), 1e-6)
        _close(out[m], ref2[m], 1e-3, 1e-4 * scale, m)


def test_fused_score_after_training():
    """Scoring with the parameters two Adam steps leave behind (every bias non-zero, weight image refreshed by the update kernel):
    MAE / Latent against the oracle evaluated on the DEVICE's parameters, rel 1e-3 -- the sequence __graft_entry__.smoke() runs."""
    from oracle import scoring as S
    FC, D, B = [80, 40, 10], 80, 512
    X, w, eps = _data(FC, D, B, 7)
    dev = _dev(FC, D); dev.set_params(V.init_params(FC, D, 0))
    for _ in range(2):
        dev.train_step(X, w, None, None, eps, 1e-3, 1.0, 0.0, 0.0)
    th = dev.get_params()
    out = {k: v.cpu().numpy() for k, v in dev.score(X, 1, eps, metrics=('MAE', 'Latent')).items()}
    xhat = S.predict(th, FC, D, X, [eps], np.float64)
    _close(out['MAE'], np.mean(np.abs(X - xhat), 1), 1e-3, 1e-6, 'MAE')
    _close(out['Latent'], S.latent_score(th, FC, D, X, np.float64), 1e-3, 1e-6, 'Latent')


def _close_inf(a, b, rtol, atol, what):
    """allclose that also wants the same +-inf entries (sums of +-FLT_MAX terms overflow in NumPy's float32 too)."""
    fa, fb = np.isfinite(a), np.isfinite(b)
    assert (fa == fb).all(), f'{what}: finite masks differ at {np.nonzero(fa != fb)[0][:5]}'
    np.testing.assert_array_equal(a[~fa], b[~fb], err_msg=what)
    if fa.any():
        _close(a[fa], b[fb], rtol, atol, what)


@pytest.mark.parametrize('case', ['zero_output', 'tiny_output', 'clipped_output', 'wild_inputs'])
def test_fused_score_exact_path(case):
    """The scoring epilogue has a range-checked fast path (one reciprocal + two logarithms per element) and an exact path
    (clip_values + the quotient form of utils.py:347-350 with NumPy's NaN -> 0, +-inf -> +-FLT_MAX rules).  These inputs
    force the exact path -- X_hat exactly 0 (dead output layer), |X_hat| < 1e-30, X_hat beyond the +-1e6 clip, huge and
    denormal-range inputs -- and must agree with NumPy's float32 result on the device's own X_hat, infinities included."""
    from atlas_vae_b200.engine import DeviceVAE
    from oracle import scoring as S
    rng = np.random.default_rng(17)
    FC, D, N = [80, 40, 20, 10], 400, 64 * 3 + 21
    flat = V.init_params(FC, D, 4).copy()
    X = rng.normal(0, 0.4, (N, D)).astype(np.float32); X[:, 300:] = 0
    n_out = FC[0] * D + D
    if case == 'zero_output':
        flat[-n_out:] = 0
    elif case == 'tiny_output':
        flat[-n_out:] = 0; flat[-D:] = rng.choice([1e-35, -1e-35, 3e-33], D).astype(np.float32)
    elif case == 'clipped_output':
        flat[-D:] += rng.choice([0, 0, 3e6, -5e7], D).astype(np.float32)
    else:
        X[::3, 5] = 2.5e6; X[1::4, 17] = -7e7; X[::5, 40] = 1e-33; X[2::7, 41] = -3e-36
    eps = rng.normal(size=(1, N, 10)).astype(np.float32)
    dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'external', 0); dev.set_params(flat)
    out = {k: v.cpu().numpy() for k, v in dev.score(X, 1, eps, metrics=('MSE', 'MAE', 'KLD', 'JSD', 'Latent'), want_xhat=True).items()}
    slim = {k: v.cpu().numpy() for k, v in dev.score(X, 1, eps, metrics=('MAE', 'KLD', 'JSD')).items()}    # instantiation without MSE
    for m in ('MAE', 'KLD', 'JSD'):
        np.testing.assert_array_equal(slim[m], out[m], err_msg=m)
    xhat = out['xhat']
    assert np.isfinite(xhat).all() and np.abs(xhat).max() <= 1e6
    if case == 'zero_output':
        assert (xhat == 0).all()
    if case == 'clipped_output':
        assert (np.abs(xhat) == 1e6).any()
    ref = S.score_metrics(X, xhat, ('MSE', 'MAE', 'KLD', 'JSD'))
    for m in ('MSE', 'MAE'):
        _close_inf(out[m], ref[m], 1e-4, 1e-6, m)
    for m in ('KLD', 'JSD'):
        fin = np.isfinite(ref[m])
        scale = max(np.abs(ref[m][fin]).max(), 1e-6) if fin.any() else 1.0
        _close_inf(out[m], ref[m], 1e-3, 1e-4 * scale, m)


@pytest.mark.parametrize('env', [{'OEVAE_SCORE_HEADS_T': '1'}, {'OEVAE_SCORE_STATIC': '1'}, {'OEVAE_SCORE_HEADS_T': '1', 'OEVAE_SCORE_STATIC': '1'}])
def test_fused_score_schedule_variants(monkeypatch, env):
    """The scoring pass hands tiles out from an atomic counter and keeps the heads with the latent index on the TMEM lanes;
    the round-robin schedule (OEVAE_SCORE_STATIC=1) and the jets-on-lanes heads (OEVAE_SCORE_HEADS_T=1, M = 64 accumulators)
    are kept as switches.  Every combination must give the default's scores: bit-equal for the schedule (a tile's arithmetic
    does not depend on which CTA runs it), and to rounding for the heads (the KLD sum is formed in another order)."""
    from atlas_vae_b200.engine import DeviceVAE
    rng = np.random.default_rng(23)
    FC, D, N = [80, 40, 20, 10], 400, 148 * 64 * 3 + 77
    X = rng.normal(0, 0.4, (N, D)).astype(np.float32); X[:, 280:] = 0
    flat = V.init_params(FC, D, 4)
    flat = (flat + rng.normal(0, 0.01, flat.shape)).astype(np.float32)     # non-zero biases in every layer
    dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 3); dev.set_params(flat)
    M = ('MSE', 'MAE', 'KLD', 'JSD', 'Latent')
    ref = {k: v.cpu().numpy() for k, v in dev.score(X, 1, None, metrics=M).items()}
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    out = {k: v.cpu().numpy() for k, v in dev.score(X, 1, None, metrics=M).items()}
    for m in M:
        if 'OEVAE_SCORE_HEADS_T' in env:
            scale = max(np.abs(ref[m]).max(), 1e-6)
            # the heads' accumulation order moves X_hat in its last bits; KLD / JSD are sums of near-cancelling log ratios
            if m in ('KLD', 'JSD'): _close(out[m], ref[m], 1e-2, 1e-3 * scale, m)
            else: _close(out[m], ref[m], 1e-3, 1e-4 * scale, m)
        else:
            np.testing.assert_array_equal(out[m], ref[m], err_msg=m)


def test_fused2_kernel_still_matches(monkeypatch):
    """The non-specialised fused2 kernel (OEVAE_FUSED_V=2) stays available as a cross-check of fused3."""
    monkeypatch.setenv('OEVAE_FUSED_V', '2')
    FC, D, B = [80, 40, 20, 10], 400, 777
    X, w, eps = _data(FC, D, B)
    flat = V.init_params(FC, D, 2)
    dev = _dev(FC, D); dev.set_params(flat)
    pj = torch.empty((4, B), device='cuda'); g = torch.empty(dev.n_params, device='cuda')
    dev.loss_and_grad(X, w, None, None, eps, 1.0, 0., 0., pj, None, g)
    monkeypatch.setenv('OEVAE_FUSED_V', '3')
    dev3 = _dev(FC, D); dev3.set_params(flat)
    pj3 = torch.empty((4, B), device='cuda'); g3 = torch.empty(dev3.n_params, device='cuda')
    dev3.loss_and_grad(X, w, None, None, eps, 1.0, 0., 0., pj3, None, g3)
    # same contractions and summation order; fused3 evaluates exp() of the Sampling layer on the SFU (rel ~1e-6)
    np.testing.assert_allclose(pj.cpu().numpy(), pj3.cpu().numpy(), rtol=2e-5, atol=1e-7)
    _close_grad(g3.cpu().numpy(), g.cpu().numpy(), 1e-4, 1e-5)
