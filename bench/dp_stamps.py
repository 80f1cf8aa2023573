"""Where the data-parallel update kernel spends its time: per-block globaltimer stamps of dp_adam_kernel (OEVAE_DP_STAMPS=1) and the
step time, per rank.   torchrun --nproc-per-node N bench/dp_stamps.py   -> one table per rank on stdout (profiles/r02_dp_stamps_*.txt)
Also measures the concurrent host->device bandwidth of all ranks (the ceiling of the e2e leg at N GPUs)."""
import os, sys, ctypes as C
os.environ['OEVAE_DP_STAMPS'] = '1'
import numpy as np, torch, torch.distributed as td
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200 import dist as Dd, _ffi
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params

rank, world = Dd.init()
local = int(os.environ.get('LOCAL_RANK', '0')); torch.cuda.set_device(local); dev = torch.device('cuda', local)
FC, D, B = [80, 40, 20, 10], 400, 5000
g = torch.Generator(device=dev); g.manual_seed(rank)
X = torch.randn((200 * B, D), device=dev, generator=g) * 0.3; w = torch.ones(B, device=dev)
v = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, local, B); v.set_params(init_params(FC, D, 0))
assert world == 1 or v.dp_connect()
step = (lambda i: v.train_step_dp(X[(i % 200) * B:(i % 200 + 1) * B], w, None, None, None, 1e-3, 1.0, 0., 0.)) if world > 1 else \
       (lambda i: v.train_step(X[(i % 200) * B:(i % 200 + 1) * B], w, None, None, None, 1e-3, 1.0, 0., 0.))
for i in range(20): step(i)
if world > 1: td.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(20, 220): step(i)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 200
rows = []
if world > 1:
    buf = np.zeros((4096, 8), np.int64)
    n = _ffi.lib().oevae_dp_stamps_host(v.h, buf.ctypes.data, buf.size)
    s = buf[:n, :6].astype(np.float64)
    t0 = s[:, 0].min()
    names = ['entry', 'dependency met (grid or early group)', 'local sum ready', 'sums pushed to the peers', 'all peer sums seen', 'exit']
    rows = ['rank %d  step %.1f us  update kernel of the last step, %d blocks (us after the first block entered)' % (rank, us, n)]
    for k, nm in enumerate(names):
        c = (s[:, k] - t0) / 1e3
        rows.append('  %-42s first %7.2f  median %7.2f  last %7.2f' % (nm, c.min(), np.median(c), c.max()))
    d = (s[:, 1:] - s[:, :-1]) / 1e3
    # the same phases by parameter group (block b owns chunk n - 1 - b of 256 parameters: layer 0 first in the flat vector, output layer last)
    nb = int(n); chunk = nb - 1 - np.arange(nb)
    P_l0, P_out = D * FC[0] + FC[0], D * FC[0] + D
    Ptot = int(_ffi.lib().oevae_n_params(v.h))
    grp = np.where((chunk + 1) * 256 <= P_l0, 0, np.where(chunk * 256 >= Ptot - P_out, 2, 1))
    for gi, gname in enumerate(['layer 0', 'chain layers (+ boundary chunks)', 'output layer']):
        sel = grp == gi
        if not sel.any(): continue
        dd = (s[sel, 1:] - s[sel, :-1]) / 1e3
        rows.append('  %-34s %3d blocks: dependency met at %6.2f..%6.2f, reduce min %.2f median %.2f max %.2f, push median %.2f max %.2f, wait for peers median %.2f max %.2f, exit at %6.2f..%6.2f'
                    % (gname, sel.sum(), (s[sel, 1].min() - t0) / 1e3, (s[sel, 1].max() - t0) / 1e3, dd[:, 1].min(), np.median(dd[:, 1]), dd[:, 1].max(), np.median(dd[:, 2]), dd[:, 2].max(),
                       np.median(dd[:, 3]), dd[:, 3].max(), (s[sel, 5].min() - t0) / 1e3, (s[sel, 5].max() - t0) / 1e3))
    rows.append('  per-block phase medians: wait for dependency %.2f, local reduce %.2f, push %.2f, wait for peers %.2f, Adam + image %.2f'
                % tuple(np.median(d, 0)))
else:
    rows = ['rank 0  step %.1f us (single rank)' % us]
# concurrent pinned host->device bandwidth of all ranks
h = torch.empty(64 << 20, dtype=torch.uint8).pin_memory(); dd = torch.empty_like(h, device=dev)
for _ in range(3): dd.copy_(h, non_blocking=True)
if world > 1: td.barrier()
torch.cuda.synchronize(); e0.record()
for _ in range(20): dd.copy_(h, non_blocking=True)
e1.record(); torch.cuda.synchronize()
rows.append('  concurrent pinned H2D bandwidth of this rank with all %d ranks copying: %.1f GB/s' % (world, 20 * h.numel() / e0.elapsed_time(e1) / 1e6))
out = [None] * world
if world > 1: td.all_gather_object(out, rows)
else: out = [rows]
if rank == 0:
    for r in out: print('\n'.join(r))
if world > 1: td.destroy_process_group()
