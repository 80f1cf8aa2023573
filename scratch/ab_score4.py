import sys, os, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
N = 148 * 64 * 64
X = torch.rand(N, D, device='cuda') * 0.3
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, 5000); dev.set_params(init_params(FC, D, 0))
M4 = ('MAE', 'KLD', 'JSD', 'Latent')
for flag in ('0', '1'):
    os.environ['OEVAE_SCORE_HEADS_T'] = flag
    dev.score(X, metrics=M4, want_finite=False); torch.cuda.synchronize()
    os.environ['OEVAE_FUSED_DBG'] = '1'
    dev.score(X, metrics=M4, want_finite=False)
    del os.environ['OEVAE_FUSED_DBG']
