import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D, B = [1024, 512, 256, 64], 400, 65536
prec = sys.argv[1] if len(sys.argv) > 1 else 'bf16x3'
X = torch.randn(B, D, device='cuda') * 0.3; w = torch.ones(B, device='cuda')
dev = DeviceVAE(FC, D, 'KLD', prec, 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
for i in range(2): dev.train_step(X, w, None, None, None, 1e-3, 1.0)
torch.cuda.synchronize()
