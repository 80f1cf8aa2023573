"""Worker phase stamps of CTA 0 for a steady-state tile of the scoring pass (OEVAE_FUSED_DBG=1 OEVAE_DBG_SKIP=k).
    OEVAE_FUSED_DBG=1 OEVAE_DBG_SKIP=10 python bench/score_stamps.py 2> stamps.txt"""
import os, sys
os.environ.setdefault('OEVAE_FUSED_DBG', '1'); os.environ.setdefault('OEVAE_DBG_SKIP', '10')
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
dev = torch.device('cuda', 0)
v = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, 0, 5000); v.set_params(init_params(FC, D, 0))
g = torch.Generator(device=dev); g.manual_seed(0)
X = torch.randn((148 * 64 * 24, D), device=dev, generator=g) * 0.3
for _ in range(2): v.score(X, metrics=('MAE', 'KLD', 'JSD', 'Latent'))
torch.cuda.synchronize()
