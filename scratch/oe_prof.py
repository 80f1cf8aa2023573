import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
from atlas_vae_b200 import _ffi
FC, D = [80, 40, 20, 10], 400
for B in (5000, 4736):
    X = torch.randn(B, D, device='cuda') * 0.3; Xo = torch.randn(B, D, device='cuda') * 0.4; w = torch.ones(B, device='cuda')
    for oe in ('KLD', 'MAE'):
        dev = DeviceVAE(FC, D, oe, 'bf16x3', 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
        for i in range(5): dev.train_step(X, w, Xo, w, None, 1e-3, 1.0, 1.0, 1.0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(20): dev.train_step(X, w, Xo, w, None, 1e-3, 1.0, 1.0, 1.0)
        e1.record(); torch.cuda.synchronize()
        print(B, oe, 'OE train_step us', 1e3 * e0.elapsed_time(e1) / 20)
        _ffi.profile(True)
        for i in range(10): dev.train_step(X, w, Xo, w, None, 1e-3, 1.0, 1.0, 1.0)
        torch.cuda.synchronize()
        for r in _ffi.profile_report(): print('   ', r)
        _ffi.profile(False)
