"""Timing (CUDA events) of the byte-bound kernels either side of the network: K0 preprocess (pT sort, slice, jet m / pT, affine scaler),
min/max + loss_mapping, K5 cut + jet-mass histogram.  Prints achieved GB/s on the algorithmic bytes."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200 import utils as U, cuts as C, _ffi
from atlas_vae_b200._ffi import ptr, check
dev = torch.device('cuda:0')
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return best
N, n_raw, n_const = int(os.environ.get('N_JETS', 1_000_000)), 200, 100
g = torch.Generator(device=dev); g.manual_seed(1)
raw = torch.randn(N, 4 * n_raw, device=dev, generator=g)
mult = torch.randint(20, 120, (N, 1), device=dev, generator=g)
raw = raw * (torch.arange(n_raw, device=dev).repeat_interleave(4)[None, :] < mult)           # zero-padded constituents
class Sc: pass
from sklearn.preprocessing import MaxAbsScaler
sc = MaxAbsScaler().fit(np.abs(np.random.default_rng(0).normal(size=(1000, 4 * n_const))).astype(np.float32))
for name, r in (('fp32', raw), ('fp16', raw.half())):
    ms = timeit(lambda: U.preprocess_device(r, n_const, 4, sc, True))
    byt = N * (4 * n_raw * r.element_size() + 4 * n_const * 4 + 8)
    print('preprocess %s raw: %.3f ms  %.1f M jets/s  %.0f GB/s on %d B/jet' % (name, ms, N / ms / 1e3, byt / ms / 1e6, byt // N), flush=True)
del raw
M = 20_000_000
score = torch.rand(M, device=dev, generator=g); mass = torch.rand(M, device=dev, generator=g) * 300
w = torch.ones(M, device=dev); jzw = torch.where(torch.rand(M, device=dev, generator=g) < 0.2, -1.0, 0.0)
keep = torch.ones(M, dtype=torch.uint8, device=dev)
edges = torch.arange(0.0, 302.5, 2.5, dtype=torch.float64, device=dev)
counts = torch.zeros((2, edges.numel() - 1), dtype=torch.int64, device=dev); wsums = torch.zeros_like(counts, dtype=torch.float64)
s = torch.cuda.current_stream().cuda_stream
ms = timeit(lambda: check(_ffi.lib().oevae_hist(ptr(mass), ptr(w), ptr(jzw), ptr(score), ptr(keep), M, 0.3, ptr(edges), edges.numel(), ptr(counts), ptr(wsums), 0, s)))
print('hist (cut + 2-class jet-mass histogram), 20 M jets: %.3f ms  %.0f GB/s on 17 B/jet' % (ms, M * 17 / ms / 1e6), flush=True)
ms = timeit(lambda: C.loss_mapping_device(score.clone(), keep))
print('min/max + loss_mapping (incl. a clone), 20 M: %.3f ms' % ms, flush=True)
print('ok')
