"""Small target for `ncu --set full`: a few cfg-2 training steps (B=5000, D=400, FC [80,40,20,10], bf16x3) and a
scoring pass over random inputs.  Only liboevae kernels are of interest (-k regex:...)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D, B = [80, 40, 20, 10], 400, 5000
NB = 8
X = torch.randn(NB * B, D, device='cuda') * 0.3
w = torch.ones(NB * B, device='cuda')
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, B)
dev.set_params(init_params(FC, D, 0))
for i in range(NB):
    dev.train_step(X[i * B:(i + 1) * B], w[i * B:(i + 1) * B], None, None, None, 1e-3, 1.0, 0., 0.)
if '--score' in sys.argv:
    dev.score(X, 1, None, metrics=('MAE', 'KLD', 'JSD', 'Latent'))
torch.cuda.synchronize()
print('ok')
