"""tcgen05.mma issue / completion rate of one CTA by operand layout (oevae_probe_issue): cycles per MMA (kind::f16, K = 16).
mode 1: A, B K-major; 2: B MN-major; 3: A and B MN-major; 4: A MN-major."""
import os, sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200 import _ffi
lib = _ffi.lib()
torch.cuda.init(); torch.zeros(1, device='cuda')
out = np.zeros(1, np.int64)
for M in (128, 64):
    for N in (16, 32, 48, 64, 80, 128):
        row = []
        for mode in (1, 2, 3, 4):
            for nw in (1,):
                nm = 240
                rc = lib.oevae_probe_issue(M, N, nw, nm, mode, out.ctypes.data_as(C.POINTER(C.c_int64)), 0)
                row.append('%6.1f' % (out[0] / nm) if rc == 0 else '   err')
        print('M=%3d N=%3d  cycles per MMA  KK %s   K.MN %s   MN.MN %s   MN.K %s' % (M, N, *row))

print('issue rate with nw warps issuing concurrently, each to its own accumulator (M=128, N=64, K-major): cycles until all have completed / MMAs per warp')
for nw in (1, 2, 3, 4):
    for N in (64, 128):
        nm = 240
        rc = lib.oevae_probe_issue(128, N, nw, nm, 1, out.ctypes.data_as(C.POINTER(C.c_int64)), 0)
        print('  nw=%d N=%3d: %6.1f cycles per MMA per warp -> %5.1f cycles per MMA overall' % (nw, N, out[0] / nm, out[0] / nm / nw))
