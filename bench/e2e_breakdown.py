"""Where the end-to-end epoch of bench.py's `e2e` leg spends its wall time: host->device bandwidth of batch-sized copies from the
registered sample, then models.train_model under cProfile (the same call bench.py times).
    python bench/e2e_breakdown.py            or            torchrun --nproc-per-node N bench/e2e_breakdown.py"""
import os, sys, time, cProfile, pstats, io
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200 import dist as Dd
from atlas_vae_b200.models import VariationalAutoEncoder, train_model
from sklearn import preprocessing

rank, world = Dd.init()
local = int(os.environ.get('LOCAL_RANK', '0')); torch.cuda.set_device(local); dev = torch.device('cuda', local)
FC, D, B, NB = [80, 40, 20, 10], 400, 5000, 400
rng = np.random.default_rng(rank)
X16 = torch.as_tensor((rng.normal(0, 30, (NB * B, D)) * (rng.uniform(size=(NB * B, D)) > 0.3)).astype(np.float16)).pin_memory().numpy()
w = torch.ones(NB * B).pin_memory().numpy()
scaler = preprocessing.MaxAbsScaler().fit(X16[:20000].astype(np.float32))
out = []

# 1. batch-sized copies from the sample, back to back on one stream
d = torch.empty((B, D), dtype=torch.float16, device=dev)
src = torch.from_numpy(X16)
for nb in (1, 4):
    for _ in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for b in range(0, NB, nb): d[:B].copy_(src[b * B:(b + 1) * B], non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    out.append('rank %d: %d copies of %.2f MB from the page-locked sample: %.1f GB/s (%.1f us per copy)' % (rank, NB // nb, B * D * 2 / 1e6, (NB // nb) * B * D * 2 / dt / 1e9, dt / (NB // nb) * 1e6))

vbk = {'constituents': X16[:B], 'weights': w[:B]}
model = VariationalAutoEncoder(FC, D, seed=0, precision='bf16x3', device=local, max_batch=B)
ts = {}
for nb in (200, 400, 200, 400):                                   # fixed cost per call vs cost per step: t(400) - t(200) = 200 steps
    bkg = {'constituents': X16[:nb * B], 'weights': w[:nb * B]}
    call = lambda: train_model(model, [(bkg, bkg)], [(vbk, vbk)], 'KLD', 1, B, 1.0, 0, 0, 1e-3, verbose=False, scaler=scaler, sharded_input=True)
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter(); call(); torch.cuda.synchronize()
        ts[nb] = (time.perf_counter() - t0) * 1e3
    out.append('rank %d: train_model on %d batches: %.2f ms (%.1f us per step)' % (rank, nb, ts[nb], ts[nb] * 1e3 / nb))
out.append('rank %d: marginal cost %.1f us per step, fixed cost per call %.2f ms' % (rank, (ts[400] - ts[200]) * 1e3 / 200, 2 * ts[200] - ts[400]))
bkg = {'constituents': X16[:200 * B], 'weights': w[:200 * B]}
call = lambda: train_model(model, [(bkg, bkg)], [(vbk, vbk)], 'KLD', 1, B, 1.0, 0, 0, 1e-3, verbose=False, scaler=scaler, sharded_input=True)
pr = cProfile.Profile(); pr.enable(); call(); torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(14)
out.append('rank %d cProfile (tottime):\n' % rank + '\n'.join(l for l in s.getvalue().splitlines() if l.strip())[:6000])
if world > 1:
    import torch.distributed as td
    allo = [None] * world; td.all_gather_object(allo, out)
    if rank == 0:
        for o in allo[:2]: print('\n'.join(o))
    td.destroy_process_group()
else:
    print('\n'.join(out))
