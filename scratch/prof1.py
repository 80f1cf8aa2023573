import sys, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D, B = [80, 40, 20, 10], 400, 5000
X = torch.randn(B, D, device='cuda') * 0.3; w = torch.ones(B, device='cuda')
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
for i in range(5): dev.train_step(X, w, None, None, None, 1e-3, 1.0, 0., 0.)
torch.cuda.synchronize()
print('ok')
