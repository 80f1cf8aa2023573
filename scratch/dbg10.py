import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
for B in (5000,):
    X = torch.randn(B, D, device='cuda') * 0.3; w = torch.ones(B, device='cuda')
    dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
    for i in range(3): dev.train_step(X, w, None, None, None, 1e-3, 1.0, 0., 0.)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(2):
        e0.record(); dev.loss_and_grad(X, w, None, None, None, 1.0, 0., 0.); e1.record(); torch.cuda.synchronize()
        print('B', B, 'loss_and_grad ms', e0.elapsed_time(e1))
    e0.record(); dev.eval_losses(X, w, None, None, None, 1.0, 0., 0.); e1.record(); torch.cuda.synchronize()
    print('B', B, 'eval ms', e0.elapsed_time(e1))
    os.environ['OEVAE_FUSED_DBG'] = '1'
    dev.loss_and_grad(X, w, None, None, None, 1.0, 0., 0.)
    dev.loss_and_grad(X, w, None, None, None, 1.0, 0., 0.)
    del os.environ['OEVAE_FUSED_DBG']
