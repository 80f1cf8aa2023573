import sys, os, torch
sys.path.insert(0, '.')
from atlas_vae_b200 import _ffi
_ffi.LIB_PATH = os.path.abspath(sys.argv[1])
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D, B = [80, 40, 20, 10], 400, 5000
X = torch.randn(200 * B, D, device='cuda') * 0.3; w = torch.ones(B, device='cuda')
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, B); dev.set_params(init_params(FC, D, 0))
def step(i): dev.train_step(X[(i % 200) * B:(i % 200 + 1) * B], w, None, None, None, 1e-3, 1.0)
for i in range(10): step(i)
torch.cuda.synchronize()
for rep in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(200): step(i)
    e1.record(); torch.cuda.synchronize()
    print(sys.argv[1], 'train_step us', 1e3 * e0.elapsed_time(e1) / 200)
