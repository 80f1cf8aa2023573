import sys, ctypes as C
sys.path.insert(0, '.')
import torch
from atlas_vae_b200 import _ffi
L = _ffi.lib()
torch.zeros(1, device='cuda')
for n_cta, same in [(1, 1), (79, 1), (148, 1), (148, 0)]:
    for bytes_, depth in [(4096, 1), (16384, 1), (32768, 1), (16384, 2), (16384, 4), (32768, 4), (16384, 8)]:
        out = (C.c_int64 * 2)()
        _ffi.check(L.oevae_probe_bulk(bytes_, depth, n_cta, same, out, 0))
        print('ctas', n_cta, 'same' if same else 'distinct', 'bytes', bytes_, 'depth', depth, 'cycles best/cold', out[0], out[1], ' B/clk/SM %.1f' % (bytes_ * depth / out[0]))
