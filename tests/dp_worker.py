"""Worker of tests/test_gpu_dp.py (run under torchrun, one rank per GPU): the peer-memory data-parallel update must give
the parameters of (a) the NCCL all-reduce path and (b) one rank stepping on the whole global batch."""
import os, sys
import numpy as np
import torch
import torch.distributed as td
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from atlas_vae_b200 import dist as Dd
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params


def main():
    rank, world = Dd.init()
    dev = torch.device('cuda', int(os.environ.get('LOCAL_RANK', '0')))
    FC, D, B, steps = [80, 40, 20, 10], 400, 1000, 6
    rng = np.random.default_rng(7)
    X = rng.normal(0, 0.5, (steps, B, D)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, (steps, B)).astype(np.float32)
    lo, hi = (B * rank) // world, (B * (rank + 1)) // world
    flat = init_params(FC, D, 1)
    mk = lambda: DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, dev.index, B)
    a, b, c = mk(), mk(), mk()
    for v in (a, b, c): v.set_params(flat)
    assert a.dp_connect(), 'peer-memory path did not connect'
    grad = torch.empty(b.n_params, device=dev)
    for s in range(steps):
        xs, ws = torch.as_tensor(X[s, lo:hi]).to(dev), torch.as_tensor(w[s, lo:hi]).to(dev)
        a.lib.oevae_set_noise_offset(a.h, lo); b.lib.oevae_set_noise_offset(b.h, lo)
        a.train_step_dp(xs, ws, None, None, None, 1e-3, 1.0, 0., 0.)
        b.loss_and_grad(xs, ws, None, None, None, 1.0, 0., 0., None, None, grad)
        td.all_reduce(grad); b.apply_update(grad, 1e-3)
        c.train_step(torch.as_tensor(X[s]).to(dev), torch.as_tensor(w[s]).to(dev), None, None, None, 1e-3, 1.0, 0., 0.)
    torch.cuda.synchronize(); td.barrier()
    pa, pb, pc = a.get_params(), b.get_params(), c.get_params()
    moved = np.abs(pc - flat).max()
    assert moved > 1e-3, 'parameters did not move'
    # replicas bit-identical across ranks
    t = torch.as_tensor(pa).to(dev); ref = t.clone(); td.broadcast(ref, 0)
    assert torch.equal(t, ref), 'replicas diverged'
    # The three paths add the per-CTA partial gradients in different orders (update kernel: 4 slices per rank, then ranks; all-reduce
    # path: 8 slices, then NCCL's tree; one rank: all tiles of the global batch), i.e. ~1e-7 relative differences in the gradient.
    # Adam's normalised update turns that into up to lr per step for the few parameters whose gradient is ~0: outlier-aware bound.
    for other, what in ((pb, 'all-reduce path'), (pc, 'one rank')):
        bad = np.abs(pa - other) > 2e-6 + 1e-4 * np.abs(other)
        assert bad.mean() < 0.03 and np.abs(pa - other).max() < steps * 1e-3 * 0.5, (what, bad.mean(), np.abs(pa - other).max())
    # ---- the outlier-exposure step (cfg-3: bkg + OoD batch, OE_type 'KLD', paired tiles) through the same update kernel
    Xo = rng.normal(0.1, 0.7, (steps, B, D)).astype(np.float32)
    wo = rng.uniform(0.5, 1.5, (steps, B)).astype(np.float32)
    a2, c2 = mk(), mk()
    for v in (a2, c2): v.set_params(flat)
    assert a2.dp_connect(), 'peer-memory path did not connect (OE engine)'
    for s in range(4):
        dv = lambda t: torch.as_tensor(t).to(dev)
        a2.lib.oevae_set_noise_offset(a2.h, lo)
        a2.train_step_dp(dv(X[s, lo:hi]), dv(w[s, lo:hi]), dv(Xo[s, lo:hi]), dv(wo[s, lo:hi]), None, 1e-3, 1.0, 1.0, 0.5)
        c2.train_step(dv(X[s]), dv(w[s]), dv(Xo[s]), dv(wo[s]), None, 1e-3, 1.0, 1.0, 0.5)
    torch.cuda.synchronize(); td.barrier()
    qa, qc = a2.get_params(), c2.get_params()
    t = torch.as_tensor(qa).to(dev); ref = t.clone(); td.broadcast(ref, 0)
    assert torch.equal(t, ref), 'OE replicas diverged'
    assert np.abs(qc - flat).max() > 1e-3
    bad = np.abs(qa - qc) > 2e-6 + 1e-4 * np.abs(qc)
    assert bad.mean() < 0.03 and np.abs(qa - qc).max() < 4 * 1e-3 * 0.5, ('OE', bad.mean(), np.abs(qa - qc).max())
    if rank == 0:
        print('dp_worker ok: world %d, max |dtheta| %.3e, dp vs nccl %.2e, dp vs 1-rank %.2e' % (world, moved, np.abs(pa - pb).max(), np.abs(pa - pc).max()))
    td.destroy_process_group()


if __name__ == '__main__':
    main()
