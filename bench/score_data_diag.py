"""Why the scoring epilogue leaves its fast path on bench.py's sample: range of the scaled inputs and of X_hat."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as Bn
from atlas_vae_b200 import synthetic as S
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
dev = torch.device('cuda', 0)
FC, D = [80, 40, 20, 10], 400
Xtr16, _, _, _ = Bn.make_files(dev, 'QCD', 200_000, 1234)
scaler = S.MaxAbs(S.maxabs_scale(Xtr16[:100_000].float()))
Xq16, _, _, _ = Bn.make_files(dev, 'QCD', 400_000, 555)
Xt16, _, _, _ = Bn.make_files(dev, 'top', 100_000, 777)
for name, x16 in (('QCD', Xq16), ('top', Xt16)):
    X = Bn.ingest(x16, scaler)
    v = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, 0, 5000); v.set_params(init_params(FC, D, 0))
    out = v.score(X, 1, None, metrics=('MAE',), want_xhat=True)
    y = out['xhat']
    ax = X.abs()
    print('%s: max |x| %.3g, fraction |x| > 1e3: %.2e, > 2.5e4: %.2e; exact zeros in x %.3f; |xhat| <= 1e-30: %.3e of entries, |x + xhat| <= 4e-30: %.3e; sum x^2 per (feature, 16 jets) >= 1e10: %.3e'
          % (name, ax.max().item(), (ax > 1e3).float().mean().item(), (ax > 2.5e4).float().mean().item(), (X == 0).float().mean().item(),
             (y.abs() <= 1e-30).float().mean().item(), ((X + y).abs() <= 4e-30).float().mean().item(),
             ((X[: (X.shape[0] // 16) * 16].view(-1, 16, D) ** 2).sum(1) >= 1e10).float().mean().item()))
    col = ax.max(0).values
    print('   columns with max |x| > 10: %d, their indices (first 20): %s' % (int((col > 10).sum()), (col > 10).nonzero().flatten()[:20].tolist()))
