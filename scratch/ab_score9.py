import sys, os, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
N = 1_000_000
X = torch.rand(N, D, device='cuda') * 0.3
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, 5000); dev.set_params(init_params(FC, D, 0))
M4 = ('MAE', 'KLD', 'JSD', 'Latent')
for i in range(20): dev.score(X, metrics=M4, want_finite=False)
torch.cuda.synchronize()
os.environ['OEVAE_FUSED_DBG'] = '1'
for n in (N, 148 * 64 * 4, N):
    dev.score(X[:n], metrics=M4, want_finite=False)
