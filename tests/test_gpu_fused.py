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
                                    ([33, 7, 5], 23, 1), ([33, 7, 5], 23, 63), ([100, 16], 130, 65), ([64, 32, 16], 64, 64 * 151 + 3)])
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
