import sys, ctypes as C
sys.path.insert(0, '.')
import torch
from atlas_vae_b200 import _ffi
L = _ffi.lib()
torch.zeros(1, device='cuda')
for M, N in [(64, 32), (64, 64), (128, 32), (128, 64), (128, 128), (64, 128)]:
    out = (C.c_int64 * 6)()
    _ffi.check(L.oevae_probe_latency(M, N, out, 0))
    print('M', M, 'N', N, 'cycles for nm=1,3,6,12,18,36:', list(out))
