import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D, B = [80, 40, 20, 10], 400, 5000
X = torch.randn(B, D, device='cuda') * 0.3; w = torch.ones(B, device='cuda')
for prec in ('fp32', 'bf16x3'):
    dev = DeviceVAE(FC, D, 'KLD', prec, 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
    for i in range(5): dev.train_step(X, w, None, None, None, 1e-3, 1.0, 0., 0.)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(100): dev.train_step(X, w, None, None, None, 1e-3, 1.0, 0., 0.)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(prec, 'host issue %.1f us/step, total %.1f us/step' % ((t1 - t0) * 1e4, (t2 - t0) * 1e4))
    # CUDA graph capture of one step
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        dev.train_step(X, w, None, None, None, 1e-3, 1.0, 0., 0.)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            dev.train_step(X, w, None, None, None, 1e-3, 1.0, 0., 0.)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(100): g.replay()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(prec, 'graph replay total %.1f us/step' % ((t2 - t0) * 1e4))
