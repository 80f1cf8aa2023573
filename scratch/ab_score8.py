import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
N = 1_000_000
X = torch.rand(N, D, device='cuda') * 0.3
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, 5000); dev.set_params(init_params(FC, D, 0))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
M4 = ('MAE', 'KLD', 'JSD', 'Latent')
def t(reps=12):
    ts = []
    for r in range(reps):
        e0.record(); dev.score(X, metrics=M4, want_finite=False); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return ts
for rnd in range(2):
    for st in ('1', '0'):
        for ht in ('0', '1'):
            os.environ['OEVAE_SCORE_STATIC'] = st; os.environ['OEVAE_SCORE_HEADS_T'] = ht
            ts = t()
            print('static', st, 'heads_t', ht, 'min %.4f median %.4f mean %.4f max %.4f ms' % (min(ts), float(np.median(ts)), float(np.mean(ts)), max(ts)), flush=True)
