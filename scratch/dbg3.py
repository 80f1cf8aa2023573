import sys, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200 import _ffi
from atlas_vae_b200._ffi import ptr, check
L = _ffi.lib()
rng = np.random.default_rng(0)
for (K, M, N) in [(16, 128, 16), (64, 128, 80), (128, 128, 128), (16, 64, 16), (64, 64, 48), (128, 64, 128), (48, 128, 96)]:
    A = torch.tensor(rng.integers(-4, 5, (K, M)).astype(np.float32)).cuda()
    B = torch.tensor(rng.integers(-4, 5, (K, N)).astype(np.float32)).cuda()
    raw = torch.full((128, N), -777.0, device='cuda')
    check(L.oevae_probe_mn(ptr(A), ptr(B), K, M, N, ptr(raw), 0, None))
    torch.cuda.synchronize()
    ref = (A.T @ B).cpu().numpy(); raw = raw.cpu().numpy()
    if M == 128:
        ok = np.array_equal(raw, ref)
        print((K, M, N), 'M=128 identity lane map:', ok)
    else:
        lanes = np.array([(i % 16) + 32 * (i // 16) for i in range(64)])
        ok = np.array_equal(raw[lanes], ref)
        print((K, M, N), 'M=64 lane map (i%16)+32*(i//16):', ok)
    if not ok:
        # find for each ref row the matching raw lane
        m = [int(np.where((raw == ref[i]).all(1))[0][0]) if (raw == ref[i]).all(1).any() else -1 for i in range(min(M, 20))]
        print('   row->lane', m, ' max abs diff', np.abs(raw[:M] - ref).max())
