"""World-size-2 checks of the data-parallel host logic on CPU (gloo): sharding of a global batch, SUM semantics of the
gradient all-reduce, the loss_mapping min/max exchange and the histogram reduce.  The per-rank
"kernel" here is the CPU oracle (tests may use it); on the GPU box the same host code drives liboevae.so over NCCL."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip('torch')
import torch.distributed as td
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from atlas_vae_b200 import dist as D
    from atlas_vae_b200.models import _BatchStager
    from oracle import vae_numpy as V
    r, w = D.init('gloo')
    assert (r, w) == (rank, world) and D.rank_world() == (rank, world)

    # ---- scoring shards: contiguous, disjoint, exhaustive
    lo, hi = D.shard_bounds(1001, rank, world)
    cover = torch.zeros(1001); cover[lo:hi] = 1
    D.all_reduce_sum(cover)
    assert bool((cover == 1).all())

    # ---- training: every global batch is split contiguously; SUM of the shard gradients == full-batch gradient
    FC, Dm, B = [12, 6, 3], 10, 37
    rng = np.random.default_rng(5)
    X = rng.normal(0, 0.6, (B, Dm)).astype(np.float32); Xo = rng.normal(0.2, 0.8, (B, Dm)).astype(np.float32)
    wt = rng.uniform(0.5, 1.5, B).astype(np.float32); eps = rng.normal(size=(B, 3)).astype(np.float32)
    flat = V.init_params(FC, Dm, 1)
    st = _BatchStager.__new__(_BatchStager)                       # bounds() only: no CUDA objects
    st.batch_size, st.n, st.rank, st.world = 16, B, rank, world
    g_sum = torch.zeros(flat.size, dtype=torch.float64); loss_sum = torch.zeros(4, dtype=torch.float64)
    rows = torch.zeros(B)
    for b in range(3):
        s0, s1, _ = st.bounds(b)
        rows[s0:s1] += 1
        if s1 > s0:
            out_, g = V.losses_and_grad(flat, FC, Dm, X[s0:s1], wt[s0:s1], Xo[s0:s1], wt[s0:s1], eps[s0:s1], OE_type='KLD',
                                        beta=0.7, lamb=0.5, margin=0.3)
            g_sum += torch.from_numpy(np.asarray(g, np.float64)); loss_sum += torch.tensor([float(np.sum(o)) for o in out_], dtype=torch.float64)
    D.all_reduce_sum(g_sum); D.all_reduce_sum(loss_sum); D.all_reduce_sum(rows)
    assert bool((rows == 1).all())                                # each jet on exactly one rank
    g_full = np.zeros(flat.size); l_full = np.zeros(4)
    for b in range(3):
        lo_, hi_ = 16 * b, min(16 * (b + 1), B)
        out_, g = V.losses_and_grad(flat, FC, Dm, X[lo_:hi_], wt[lo_:hi_], Xo[lo_:hi_], wt[lo_:hi_], eps[lo_:hi_], OE_type='KLD',
                                    beta=0.7, lamb=0.5, margin=0.3)
        g_full += g; l_full += [float(np.sum(o)) for o in out_]
    np.testing.assert_allclose(g_sum.numpy(), g_full, rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(loss_sum.numpy(), l_full, rtol=1e-12)

    # ---- loss_mapping predicates are global: min / max over all shards
    score = torch.from_numpy(rng.normal(size=64).astype(np.float32))      # same stream on both ranks
    lo_t = score[rank * 32:(rank + 1) * 32].min().double().reshape(1); hi_t = score[rank * 32:(rank + 1) * 32].max().double().reshape(1)
    D.all_reduce_minmax(lo_t, hi_t)
    assert float(lo_t) == float(score.min()) and float(hi_t) == float(score.max())

    # ---- histogram reduce: integer counts are order-independent => bit-exact
    m = rng.uniform(0, 300, 5000).astype(np.float32)
    edges = np.append(np.arange(0, 300, 2.5), 300.0)
    a, b_ = D.shard_bounds(m.size, rank, world)
    counts = torch.from_numpy(np.histogram(m[a:b_], edges)[0].astype(np.int64))
    D.all_reduce_sum(counts)
    assert np.array_equal(counts.numpy(), np.histogram(m, edges)[0])
    D.barrier()
    out.put(rank)
    td.destroy_process_group()


def test_world2_gloo_data_parallel_logic():
    ctx = mp.get_context('spawn')
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs: p.start()
    for p in procs: p.join(180)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert sorted(out.get(timeout=5) for _ in range(2)) == [0, 1]
