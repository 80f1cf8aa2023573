import sys, os, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
N = 148 * 64 * 100
X = torch.rand(N, D, device='cuda') * 0.3
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, 5000); dev.set_params(init_params(FC, D, 0))
M4 = ('MAE', 'KLD', 'JSD', 'Latent')
os.environ['OEVAE_SCORE_HEADS_T'] = '1'
for skip in ('0', '40', '90'):
    dev.score(X, metrics=M4, want_finite=False); torch.cuda.synchronize()
    os.environ['OEVAE_FUSED_DBG'] = '1'; os.environ['OEVAE_DBG_SKIP'] = skip
    print('skip', skip, file=sys.stderr, flush=True)
    dev.score(X, metrics=M4, want_finite=False)
    del os.environ['OEVAE_FUSED_DBG']
