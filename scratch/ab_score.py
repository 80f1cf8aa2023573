import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
N = 1_000_000
X = torch.rand(N, D, device='cuda') * 0.3
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, 5000); dev.set_params(init_params(FC, D, 0))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def t(metrics, reps=5):
    best = 1e9
    for r in range(reps):
        e0.record(); o = dev.score(X, metrics=metrics, want_finite=False); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, o
for rnd in range(2):
    for flag in ('0', '1'):
        os.environ['OEVAE_SCORE_HEADS_T'] = flag
        for m in (('MAE', 'KLD', 'JSD', 'Latent'), ('MAE', 'Latent')):
            ms, o = t(m)
            print('heads_t', flag, m, 'ms', round(ms, 4), 'Mjets/s', round(N / ms / 1e3, 1), 'Latent sum', float(o['Latent'].double().sum()), flush=True)
os.environ['OEVAE_FUSED_DBG'] = '1'
dev.score(X[:148 * 64 * 4], metrics=('MAE', 'KLD', 'JSD', 'Latent'), want_finite=False)
