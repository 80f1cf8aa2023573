import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
from atlas_vae_b200 import _ffi
FC, D, B = [1024, 512, 256, 64], 400, 65536
X = torch.randn(B, D, device='cuda') * 0.3; w = torch.ones(B, device='cuda')
for prec in ('bf16x3', 'bf16', 'tf32'):
    try:
        dev = DeviceVAE(FC, D, 'KLD', prec, 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
    except Exception as e:
        print(prec, 'unavailable', e); continue
    for i in range(3): dev.train_step(X, w, None, None, None, 1e-3, 1.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(10): dev.train_step(X, w, None, None, None, 1e-3, 1.0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flops = 6.0 * B * sum(a * b for a, b in zip([D] + FC[:-1] + [FC[-1]] + FC[-2::-1], FC[:-1] + [2 * FC[-1]] + FC[-2::-1] + [D]))
    print(prec, 'cfg5 train_step ms', ms, 'jets/s', B / ms * 1e3, 'TFLOP/s (algorithmic, 1 pass)', flops / ms / 1e9)
    _ffi.profile(True)
    for i in range(3): dev.train_step(X, w, None, None, None, 1e-3, 1.0)
    torch.cuda.synchronize()
    rep = sorted(_ffi.profile_report(), key=lambda r: -r[2])
    for r in rep[:8]: print('   ', r)
    _ffi.profile(False)
