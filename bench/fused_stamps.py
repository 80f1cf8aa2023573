"""Phase stamps of CTA 0 of the fused training kernel (OEVAE_FUSED_DBG=1; synchronises every launch): cycles since the first stamp.
    OEVAE_FUSED_DBG=1 python bench/fused_stamps.py 2> stamps.txt"""
import os, sys
os.environ.setdefault('OEVAE_FUSED_DBG', '1')
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D, B = [80, 40, 20, 10], 400, 5000
dev = torch.device('cuda', 0)
v = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, 0, B); v.set_params(init_params(FC, D, 0))
g = torch.Generator(device=dev); g.manual_seed(0)
X = torch.randn((8 * B, D), device=dev, generator=g) * 0.3; w = torch.ones(B, device=dev)
for i in range(8): v.train_step(X[i * B:(i + 1) * B], w, None, None, None, 1e-3, 1.0, 0., 0.)
torch.cuda.synchronize()
