import sys, ctypes as C
sys.path.insert(0, '.')
import torch
from atlas_vae_b200 import _ffi
L = _ffi.lib()
torch.zeros(1, device='cuda')
for M, N in [(128, 64), (128, 128)]:
    for mode in (0, 1):
        for nw in (1, 2, 4):
            row = []
            for nm in (12, 48):
                out = C.c_int64()
                _ffi.check(L.oevae_probe_issue(M, N, nw, nm, mode, C.byref(out), 0))
                row.append(out.value)
            print('M', M, 'N', N, 'mode', mode, 'warps', nw, 'cycles nm=12,48 per warp:', row, ' per-MMA(total) %.1f' % ((row[1] - row[0]) / 36 / nw))
