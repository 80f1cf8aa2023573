import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200 import utils as U
dev = torch.device('cuda:0')
N, n_raw, n_const = 200_000, 200, 100
g = torch.Generator(device=dev); g.manual_seed(1)
raw = torch.randn(N, 4 * n_raw, device=dev, generator=g)
mult = torch.randint(20, 120, (N, 1), device=dev, generator=g)
raw = raw * (torch.arange(n_raw, device=dev).repeat_interleave(4)[None, :] < mult)
from sklearn.preprocessing import MaxAbsScaler
sc = MaxAbsScaler().fit(np.abs(np.random.default_rng(0).normal(size=(1000, 4 * n_const))).astype(np.float32))
for i in range(3): U.preprocess_device(raw, n_const, 4, sc, True)
torch.cuda.synchronize(); print('ok')
