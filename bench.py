#!/usr/bin/env python
"""bench.py -- OE-VAE training throughput (jets/s, fwd+bwd+Adam) on N B200s; BASELINE.json config 2 is the headline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp32|bf16x3|bf16]

A step = one pass of the hot path (get_losses + gradient + clip + Adam, models.py:180-186) over one batch of 5000 synthetic QCD jets
(n_constituents=100 x 4-vectors, FC_layers [80,40,20,10]).  The workload does NOT depend on --steps: the resident epoch is always
1M jets (200 batches), the scoring shard always 2.5M jets per GPU.  Prints ONE JSON line.
  value        device-timed whole-job jets/s with the 1M-jet epoch resident in HBM (every step reads a different 8 MB batch of the
               1.6 GB array: inputs larger than L2)
  e2e          the same metric through the public API: models.train_model on a HOST sample (float16 constituents as the reference's
               files store them + the MaxAbs scaler), every step's batch copied host->device from page-locked memory and the
               running loss sums copied back device->host after every step, all inside the timed region
  roofline     dominant kernel (per-kernel CUDA-event pass), algorithmic bytes of SURVEY 8(d) vs MEASURED_PEAKS.json
  scoring      cfg-4: anomaly-scoring pass (predict + Latent/MAE/KLD/JSD + loss_mapping + cut + jet-mass histogram, reduced across
               ranks) over 2.5M jets per GPU
  oe / wide    cfg-3 (paired OoD batch, OE_type KLD) and cfg-5 (FC_layers [1024,512,256,64], batch 65536) steps, at every N
  strong       strong scaling: global batch fixed at 5000 and at 40000, split over the ranks
  small_kernels  K0 preprocessing, histogram, min/max, loss map, Adam: each with its own roofline
  dp_parity    (N > 1) 3-step self-check before timing: exchanged gradient sum vs oevae_loss_and_grad of the global batch on rank 0,
               replicas bit-identical
  cpu_baseline the reference step restated op-for-op in torch-CPU eager fp32 (oracle/torch_twin.py) on the host cores of this box
--impl reference times only that CPU restatement (TensorFlow 2.1 cannot be installed here).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FC, N_CONST, N_DIMS, BATCH, N_JETS = [80, 40, 20, 10], 100, 4, 5000, 1_000_000
N_SCORE = 2_500_000                                 # cfg-4: 20M test jets over 8 GPUs = 2.5M per GPU
D = N_CONST * N_DIMS
ALG_BYTES_PER_JET = D * 4 + 4                       # X row + weight (1604 B)
SC_ALG = D * 4 + 4 + 4 + 1 + 4 * 4                  # X row, m, w, label, 4 scores = 1625 B
BETA, LR = 1.0, 1e-3
METRIC, UNIT = 'train jets/s (fwd+bwd+Adam)', 'jets/s'
# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures (profiles/): a profile of an
# earlier build of the same kernel, NOT measured in this run
PROFILE_TRAFFIC = {'fused_train': {'bytes_per_launch': 8575744 + 11264, 'from_profile': 'profiles/r02b_ncu_fused3_adam_summary.txt (fused3_kernel<0,0>, B=5000: dram__bytes_read + dram__bytes_write)'},
                   'fused_score': {'bytes_per_jet': (64572160 + 1385728) / 40000.0, 'from_profile': 'profiles/r01_ncu_score_fused3_details.txt (40,000 jets)'}}


def workload_text(n_jets):
    return ('cfg2: OE-VAE train step, n_constituents=%d x 4-vectors (D=%d), FC_layers %s, batch %d, %s synthetic QCD jets resident '
            '(%d distinct batches), beta=%g, no OoD term' % (N_CONST, D, FC, BATCH, '{:,}'.format(n_jets), n_jets // BATCH, BETA))


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        return float(p['hbm_gbs']), float(p.get('bf16_tflops_sustained', 1400.0)), 'measured'
    except Exception:
        return 6650.0, 1400.0, 'fallback'


class ClockSampler:
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits', '-lms', '100'],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.p.terminate(); self.p.wait()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.f.read().splitlines():
            c = [v.strip() for v in line.split(',')]
            if len(c) < 6: continue
            try: sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError: continue
            for nme, v in zip(names, c[2:6]):
                if v.lower().startswith('active'): reasons.add(nme)
        os.unlink(self.f.name)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None, 'reasons': sorted(reasons), 'samples': len(sm)}


def cpu_reference_steps(steps, warmup, B=BATCH, budget_s=60.0):
    """The reference CPU path (oracle/torch_twin.py, kind='port'): returns (jets/s, cores, seconds, steps)."""
    import numpy as np
    import torch
    from oracle import torch_twin as T, vae_numpy as V
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    rng = np.random.default_rng(0)
    vae = T.TorchVAE(V.init_params(FC, D, 0), FC, D, torch.float32)
    n_b = 4
    Xs = [torch.tensor(rng.normal(0, 0.3, (B, D)).astype(np.float32)) for _ in range(n_b)]
    w = torch.ones(B); eps = torch.tensor(rng.normal(size=(B, FC[-1])).astype(np.float32))
    dt = 0.0; done = 0
    warmup = max(warmup, 5)                          # the first torch-CPU steps pay thread-pool / allocator start-up
    for s in range(warmup + steps):
        if dt > budget_s: break
        X = Xs[s % n_b]
        t0 = time.perf_counter()
        # no OoD sample in config 2: the OE term is evaluated on the batch itself with lamb=0, as vae.py does when lamb==0
        T.train_step(vae, X, w, X, w, eps, 'KLD', BETA, 0.0, 0.0, LR)
        if s >= warmup: dt += time.perf_counter() - t0; done += 1
    return done * B / dt, cores, dt, done


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    steps = args.steps
    jps, cores, dt, steps = cpu_reference_steps(steps, args.warmup)
    line = {'impl': 'reference', 'metric': METRIC, 'value': jps, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': steps, 'warmup': args.warmup,
            'ms_per_step': 1e3 * dt / steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': workload_text(N_JETS)},
            'cpu_baseline': {'value': jps, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                             'sample': '%d steps of %d jets: models.py:108-132,181-186 restated op-for-op in torch-CPU eager fp32 (TensorFlow 2.1 is not installable here)' % (steps, BATCH)},
            'e2e': {'value': jps, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line))


def make_files(device, kind, n_jets, seed, n_raw=N_CONST):
    """Synthetic jets as the reference's files hold them (tools/merging.py:31): pT-sorted constituents, first N_CONST kept,
    float16 [n, 4 n_const], plus jet mass / pT (utils.py:300-308) and the JZW label."""
    import torch
    from atlas_vae_b200 import synthetic as S, utils as U
    X16 = torch.empty((n_jets, D), dtype=torch.float16, device=device)
    m = torch.empty(n_jets, dtype=torch.float32, device=device); pt = torch.empty_like(m)
    jzw = torch.empty(n_jets, dtype=torch.float32, device=device)
    chunk = 250_000
    for lo in range(0, n_jets, chunk):
        n = min(chunk, n_jets - lo)
        raw, jz = S.make_jets(kind, n, n_raw, seed=seed + lo, device=device)
        x, p, mm = U.preprocess_device(raw, N_CONST, N_DIMS, None, True)
        X16[lo:lo + n] = x.to(torch.float16); m[lo:lo + n] = mm; pt[lo:lo + n] = p; jzw[lo:lo + n] = jz
        del raw, x
    return X16, m, pt, jzw


def ingest(X16, scaler):
    """float16 file rows -> scaled float32 model inputs on the device (oevae_ingest_f16)."""
    import torch
    from atlas_vae_b200 import _ffi
    from atlas_vae_b200._ffi import ptr, check
    dev = X16.device
    s_t = torch.as_tensor(scaler.scale_).to(dev)
    X = torch.empty(X16.shape, dtype=torch.float32, device=dev)
    check(_ffi.lib().oevae_ingest_f16(ptr(X16), ptr(X), X16.shape[0], X16.shape[1], None, ptr(s_t), dev.index, torch.cuda.current_stream().cuda_stream))
    return X


def run_ours(args):
    real_stdout = os.dup(1)                                       # native libraries (NCCL's version banner) write to fd 1:
    os.dup2(2, 1)                                                 # keep stdout for the one JSON line
    import numpy as np
    import torch
    import torch.distributed as td
    from atlas_vae_b200 import _ffi, dist as Dd, synthetic as S, utils as U, cuts as Cc
    from atlas_vae_b200._ffi import ptr, check
    from atlas_vae_b200.engine import DeviceVAE
    from atlas_vae_b200.models import init_params, VariationalAutoEncoder, train_model
    local = int(os.environ.get('LOCAL_RANK', '0'))
    numa = Dd.bind_to_gpu_numa(local)                             # host buffers of this rank are then first-touched on the GPU's node
    rank, world = Dd.init()
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    K, W = args.steps, max(args.warmup, 3)
    hbm_peak, tc_peak, peak_src = peaks()
    stream = lambda: torch.cuda.current_stream().cuda_stream
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def sync():
        if world > 1: td.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1: return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, n_warm, n):
        """ms per call of fn(i): n_warm untimed, n timed between barrier + synchronize, max over ranks."""
        for i in range(n_warm): fn(i)
        sync()
        e0.record()
        for i in range(n_warm, n_warm + n): fn(i)
        e1.record()
        sync()
        return max_over_ranks(e0.elapsed_time(e1)) / n

    # ---- workload: every rank owns its own 1M-jet epoch (weak scaling); the scaler is MaxAbs fitted on the first 100k jets
    n_jets = N_JETS
    n_batches = n_jets // BATCH
    X16, _, _, _ = make_files(dev, 'QCD', n_jets, 1234 + 7919 * rank)
    scaler = S.MaxAbs(S.maxabs_scale(X16[:100_000].float()))
    X = ingest(X16, scaler)
    wts = torch.ones(n_jets, dtype=torch.float32, device=dev)
    vae = DeviceVAE(FC, D, 'KLD', args.precision, 'fixed', 0, local, BATCH)
    vae.set_params(init_params(FC, D, 0))
    sums = torch.zeros(4, dtype=torch.float64, device=dev)
    grad = torch.empty(vae.n_params, dtype=torch.float32, device=dev)

    # data parallel: the update kernel exchanges the per-rank gradient sums over NVLink peer memory (dp_kernels.cuh);
    # if that path does not apply, SUM all-reduce of the flat gradient over NCCL
    peer = world > 1 and not os.environ.get('OEVAE_DP_NCCL') and vae.dp_connect()

    def make_step(v, Xs, ws, B, is_peer, g, Xo=None, lamb=0.0, margin=0.0):
        nb = Xs.shape[0] // B
        nbo = Xo.shape[0] // B if Xo is not None else 1
        def step(i):
            b = i % nb
            xb, wb = Xs[b * B:(b + 1) * B], ws[:B]
            xo = wo = None
            if Xo is not None:
                o = i % nbo
                xo, wo = Xo[o * B:(o + 1) * B], ws[:B]
            if world == 1:
                v.train_step(xb, wb, xo, wo, None, LR, BETA, lamb, margin, None, sums)
            elif is_peer:
                v.train_step_dp(xb, wb, xo, wo, None, LR, BETA, lamb, margin, None, sums)
            else:
                v.loss_and_grad(xb, wb, xo, wo, None, BETA, lamb, margin, None, sums, g)
                td.all_reduce(g, op=td.ReduceOp.SUM)
                v.apply_update(g, LR)
        return step

    # ---- data-parallel self-check (N > 1): the sum the update kernel exchanges vs the gradient of the global batch on one rank
    dp_parity = None
    if world > 1:
        Bg = BATCH
        gen = torch.Generator(device=dev); gen.manual_seed(4242)
        pick = torch.randint(0, 100_000, (3, Bg), generator=gen, device=dev)
        X16g, _, _, _ = make_files(dev, 'QCD', 100_000, 999)                 # the same rows on every rank
        Xg = ingest(X16g, scaler)
        # the weak-scaling scaler differs per rank; parity needs identical inputs: scale with rank 0's
        sc0 = torch.as_tensor(scaler.scale_).to(dev); td.broadcast(sc0, 0)
        Xg = ingest(X16g, S.MaxAbs(sc0.cpu().numpy()))
        a = DeviceVAE(FC, D, 'KLD', args.precision, 'fixed', 0, local, BATCH)
        r = DeviceVAE(FC, D, 'KLD', args.precision, 'fixed', 0, local, BATCH)
        a.set_params(init_params(FC, D, 0))
        ok_conn = (not os.environ.get('OEVAE_DP_NCCL')) and a.dp_connect()
        rel_l2, max_rel = [], []
        if ok_conn:
            lo, hi = (Bg * rank) // world, (Bg * (rank + 1)) // world
            for s_ in range(3):
                xg = Xg[pick[s_]].contiguous(); wg = torch.ones(Bg, device=dev)
                r.set_params(a.get_params())
                gref = r.loss_and_grad(xg, wg, None, None, None, BETA, 0., 0.).cpu().numpy().astype(np.float64)
                a.set_noise_offset(lo)
                a.train_step_dp(xg[lo:hi].contiguous(), wg[lo:hi].contiguous(), None, None, None, LR, BETA, 0., 0.)
                gdp = a.dp_last_grad().astype(np.float64)
                rel_l2.append(float(np.linalg.norm(gdp - gref) / max(np.linalg.norm(gref), 1e-30)))
                max_rel.append(float(np.max(np.abs(gdp - gref)) / max(np.max(np.abs(gref)), 1e-30)))
            a.set_noise_offset(0)
            th = torch.as_tensor(a.get_params()).to(dev); th0 = th.clone(); td.broadcast(th0, 0)
            same = torch.tensor([1.0 if torch.equal(th, th0) else 0.0], device=dev)
            td.all_reduce(same, op=td.ReduceOp.MIN)
            worst = torch.tensor([max(rel_l2)], dtype=torch.float64, device=dev); td.all_reduce(worst, op=td.ReduceOp.MAX)
            dp_parity = {'steps': 3, 'global_batch': Bg, 'grad_rel_l2_max_over_ranks': float(worst.item()), 'grad_rel_l2': rel_l2, 'grad_max_abs_over_max': max_rel,
                         'bound_rel_l2': 1e-5, 'replicas_bit_identical': bool(same.item() == 1.0),
                         'ok': bool(worst.item() <= 1e-5 and same.item() == 1.0),
                         'what': 'SUM over ranks of the gradient inside dp_adam_kernel (before clip) vs oevae_loss_and_grad of the global batch on each rank'}
        else:
            dp_parity = {'ok': None, 'what': 'peer-memory update not connected (NCCL all-reduce path)'}
        del a, r, Xg, X16g

    # ---- headline: device-resident cfg-2 steps
    step = make_step(vae, X, wts, BATCH, peer, grad)
    for i in range(W): step(i)
    sync()
    clocks = ClockSampler(local) if rank == 0 else None
    n0 = _ffi.launch_count()
    e0.record()
    for i in range(W, W + K): step(i)
    e1.record()
    sync()
    launches = _ffi.launch_count() - n0
    ms = max_over_ranks(e0.elapsed_time(e1))
    value = world * K * BATCH / (ms * 1e-3)
    loss_mean = float(sums[3].item()) / ((K + W) * BATCH)

    # ---- per-kernel pass (CUDA events around every launch) for the roofline leg
    _ffi.profile(True)
    kp = max(20, min(K, 50))
    for i in range(kp): step(W + K + i)
    torch.cuda.synchronize()
    rows = _ffi.profile_report()
    _ffi.profile(False)
    tot = sum(r_[2] for r_ in rows) or 1.0
    kernels = [{'name': r_[0], 'launches_per_step': r_[1] / kp, 'us_per_launch': 1e3 * r_[2] / r_[1], 'share': r_[2] / tot} for r_ in rows]
    top = kernels[0] if kernels else None
    roof = None
    if top:
        # algorithmic bytes of one launch of the dominant kernel = (bytes/jet of SURVEY 8.0) x jets it processes
        alg = ALG_BYTES_PER_JET * BATCH / max(top['launches_per_step'], 1e-9)
        ach = alg / (top['us_per_launch'] * 1e-6) / 1e9
        pt_ = PROFILE_TRAFFIC.get(top['name'])
        roof = {'bound': 'hbm', 'kernel': top['name'], 'achieved': ach, 'peak': hbm_peak, 'unit': 'GB/s', 'frac': ach / hbm_peak,
                'traffic': pt_['bytes_per_launch'] if pt_ and 'bytes_per_launch' in pt_ else None, 'traffic_from_profile': pt_['from_profile'] if pt_ else None,
                'peak_source': peak_src, 'alg_bytes_per_launch': alg, 'us_per_launch': top['us_per_launch'], 'share_of_step': top['share'],
                'step': {'achieved': ALG_BYTES_PER_JET * BATCH / (ms * 1e-3 / K) / 1e9, 'frac': ALG_BYTES_PER_JET * BATCH / (ms * 1e-3 / K) / 1e9 / hbm_peak}}
    P = vae.n_params
    small = {}
    for kk in kernels:
        if kk['name'] in ('adam', 'dp_adam'):
            n_parts = min((BATCH + 63) // 64, 148)
            alg_b = 28 * P                                             # theta, m, v read + written, g read
            small[kk['name']] = {'us_per_launch': kk['us_per_launch'], 'alg_bytes_per_launch': alg_b,
                                 'roofline': {'bound': 'hbm', 'achieved': alg_b / (kk['us_per_launch'] * 1e-6) / 1e9, 'peak': hbm_peak, 'unit': 'GB/s',
                                              'frac': alg_b / (kk['us_per_launch'] * 1e-6) / 1e9 / hbm_peak},
                                 'l2_bytes_read_partials': n_parts * P * 4,
                                 'note': 'also sums %d per-CTA partial-gradient slices from L2 and rewrites the bf16 weight image' % n_parts}

    # ---- strong scaling: the GLOBAL batch is fixed (5000 and 40000 jets) and split over the ranks
    strong = {}
    for Bg in (5000, 40000):
        Bl = Bg // world
        if Bl < 1: continue
        vs = DeviceVAE(FC, D, 'KLD', args.precision, 'fixed', 0, local, Bl)
        vs.set_params(init_params(FC, D, 0))
        p_s = world > 1 and not os.environ.get('OEVAE_DP_NCCL') and vs.dp_connect()
        st_ms = timed(make_step(vs, X, wts, Bl, p_s, grad), 5, 50)
        strong['global_batch_%d' % Bg] = {'value': Bg / (st_ms * 1e-3), 'unit': UNIT, 'ms_per_step': st_ms, 'per_gpu_batch': Bl, 'scaling': 'strong',
                                          'roofline_frac_hbm': ALG_BYTES_PER_JET * Bg / (st_ms * 1e-3) / 1e9 / (hbm_peak * world)}
        del vs

    # ---- cfg-4 anomaly-scoring leg: vae.py:163-167 predict + loss_function(Latent, MAE, KLD, JSD) + filtering flag, loss_mapping of
    # the cut metric (global min/max all-reduced), then the apply_cut jet-mass histogram (plots.py:259,697-715) of the jets above the
    # cut, counts SUM-reduced across ranks inside the timed region.  2.5M jets per GPU (16M QCD + 4M top over 8 shards); every pass
    # reads the whole 4 GB shard.
    n_top = N_SCORE // 5
    Xq16, mq, _, jq = make_files(dev, 'QCD', N_SCORE - n_top, 555 + 31 * rank)
    Xt16, mt, _, jt = make_files(dev, 'top', n_top, 777 + 31 * rank)
    Xs = torch.empty((N_SCORE, D), dtype=torch.float32, device=dev)
    Xs[:N_SCORE - n_top] = ingest(Xq16, scaler); Xs[N_SCORE - n_top:] = ingest(Xt16, scaler)
    del Xq16, Xt16
    mass = torch.cat([mq, mt]); jzw = torch.cat([jq, jt]); ws_ = torch.ones(N_SCORE, dtype=torch.float32, device=dev)
    hi_m = float(mass.max().item())
    if world > 1:
        t_ = torch.tensor([hi_m], dtype=torch.float64, device=dev); td.all_reduce(t_, op=td.ReduceOp.MAX); hi_m = float(t_.item())
    edges = torch.as_tensor(U.get_idx(np.float32(hi_m), bin_size=2.5, min_val=0, integer=False, tuples=False)).to(dev)
    counts = torch.zeros((2, edges.numel() - 1), dtype=torch.int64, device=dev); wsums = torch.zeros_like(counts, dtype=torch.float64)
    def score_pass(i=0):
        out = vae.score(Xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
        Cc.loss_mapping_device(out['MAE'], out['finite'])
        counts.zero_(); wsums.zero_()
        check(_ffi.lib().oevae_hist(ptr(mass), ptr(ws_), ptr(jzw), ptr(out['MAE']), ptr(out['finite']), N_SCORE, 0.3,
                                    ptr(edges), edges.numel(), 1.0, 0, ptr(counts), ptr(wsums), local, stream()))
        if world > 1:
            td.all_reduce(counts, op=td.ReduceOp.SUM); td.all_reduce(wsums, op=td.ReduceOp.SUM)
    sc_ms = timed(score_pass, 2, 5)
    _ffi.profile(True)                                            # a separate pass for the kernel shares (not the timed one)
    score_pass(); torch.cuda.synchronize()
    srows = _ffi.profile_report(); _ffi.profile(False)
    stot = sum(r_[2] for r_ in srows) or 1.0
    sc_val = world * N_SCORE / (sc_ms * 1e-3)
    sc_top = srows[0] if srows else None
    scoring = {'metric': 'anomaly-score jets/s (predict + Latent/MAE/KLD/JSD + loss_mapping + cut + jet-mass histogram reduced over ranks)', 'value': sc_val, 'unit': UNIT,
               'n_jets_per_gpu': N_SCORE, 'n_jets_total': world * N_SCORE, 'ms_per_pass': sc_ms, 'hist_bins': int(edges.numel() - 1),
               'hist_entries': int(counts.sum().item()),
               'roofline': {'bound': 'hbm', 'achieved': N_SCORE * SC_ALG / (sc_ms * 1e-3) / 1e9, 'peak': hbm_peak, 'unit': 'GB/s',
                            'frac': N_SCORE * SC_ALG / (sc_ms * 1e-3) / 1e9 / hbm_peak, 'alg_bytes_per_jet': SC_ALG,
                            'kernel': sc_top[0] if sc_top else None,
                            'kernel_frac': (N_SCORE * SC_ALG / (1e-3 * sc_top[2] / sc_top[1]) / 1e9 / hbm_peak) if sc_top else None,
                            'traffic': PROFILE_TRAFFIC['fused_score']['bytes_per_jet'] * N_SCORE, 'traffic_from_profile': PROFILE_TRAFFIC['fused_score']['from_profile']},
               'kernels': [{'name': r_[0], 'launches_per_pass': r_[1], 'us_per_launch': 1e3 * r_[2] / r_[1], 'share': r_[2] / stot} for r_ in srows[:6]]}
    if os.environ.get('BENCH_DEBUG_SCORE'):                       # diagnostic: the scoring kernel alone, as bench/score_sweep.py times it
        for tag, xs in (('Xs', Xs), ('Xs.clone()', Xs.clone())):
            for _ in range(2): vae.score(xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
            torch.cuda.synchronize(); e0.record()
            for _ in range(4): vae.score(xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
            e1.record(); torch.cuda.synchronize()
            print('[bench debug] vae.score(%s) alone: %.3f ms per pass' % (tag, e0.elapsed_time(e1) / 4), file=sys.stderr)
        v2 = DeviceVAE(FC, D, 'KLD', args.precision, 'fixed', 0, local, BATCH); v2.set_params(vae.get_params())
        for _ in range(2): v2.score(Xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
        torch.cuda.synchronize(); e0.record()
        for _ in range(4): v2.score(Xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
        e1.record(); torch.cuda.synchronize()
        print('[bench debug] fresh handle, same weights: %.3f ms per pass' % (e0.elapsed_time(e1) / 4), file=sys.stderr)
        from atlas_vae_b200.models import init_params as _ip
        v2.set_params(_ip(FC, D, 0))
        for tag, xs in (('init weights, Xs', Xs), ('init weights, randn', torch.randn_like(Xs) * 0.3)):
            for _ in range(2): v2.score(xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
            torch.cuda.synchronize(); e0.record()
            for _ in range(4): v2.score(xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
            e1.record(); torch.cuda.synchronize()
            print('[bench debug] %s: %.3f ms per pass' % (tag, e0.elapsed_time(e1) / 4), file=sys.stderr)
        v2.set_params(vae.get_params())
        xs = torch.randn_like(Xs) * 0.3
        for _ in range(2): v2.score(xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
        torch.cuda.synchronize(); e0.record()
        for _ in range(4): v2.score(xs, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
        e1.record(); torch.cuda.synchronize()
        print('[bench debug] trained weights, randn: %.3f ms per pass' % (e0.elapsed_time(e1) / 4), file=sys.stderr)
    clk = clocks.stop() if clocks else None

    # ---- small kernels, each with its own roofline (device-timed on this rank's data)
    def one(name, fn, n_units, alg_bytes, n=10):
        for _ in range(2): fn()
        torch.cuda.synchronize(); e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        t_ms = e0.elapsed_time(e1) / n
        small[name] = {'ms_per_launch': t_ms, 'units_per_s': n_units / (t_ms * 1e-3), 'alg_bytes_per_unit': alg_bytes,
                       'roofline': {'bound': 'hbm', 'achieved': n_units * alg_bytes / (t_ms * 1e-3) / 1e9, 'peak': hbm_peak, 'unit': 'GB/s',
                                    'frac': n_units * alg_bytes / (t_ms * 1e-3) / 1e9 / hbm_peak}}
    if not args.no_extra:
        n_k0 = 400_000
        raw, _ = S.make_jets('QCD', n_k0, 200, seed=31 + rank, device=dev)             # 200 stored constituents -> keep 100 (SURVEY: 4,808 B per jet)
        sc_t = torch.as_tensor(scaler.scale_).to(dev)
        xk = torch.empty((n_k0, D), dtype=torch.float32, device=dev); pk = torch.empty(n_k0, device=dev); mk_ = torch.empty(n_k0, device=dev)
        one('k0_preprocess', lambda: check(_ffi.lib().oevae_preprocess(ptr(raw), 0, n_k0, 200, N_CONST, N_DIMS, None, ptr(sc_t), 1, ptr(xk), ptr(pk), ptr(mk_), local, stream())),
            n_k0, 200 * 16 + D * 4 + 8)
        del raw, xk
        sc_buf = torch.rand(N_SCORE, device=dev)
        fin = torch.ones(N_SCORE, dtype=torch.uint8, device=dev)
        one('hist', lambda: check(_ffi.lib().oevae_hist(ptr(mass), ptr(ws_), ptr(jzw), ptr(sc_buf), ptr(fin), N_SCORE, 0.3, ptr(edges), edges.numel(), 1.0, 0, ptr(counts), ptr(wsums), local, stream())),
            N_SCORE, 17)
        mm_ = torch.tensor([np.inf, -np.inf, 0.0], dtype=torch.float64, device=dev)
        one('minmax', lambda: check(_ffi.lib().oevae_minmax(ptr(sc_buf), ptr(fin), N_SCORE, ptr(mm_), local, stream())), N_SCORE, 5)
        one('loss_map', lambda: check(_ffi.lib().oevae_loss_map(ptr(sc_buf), N_SCORE, 2, local, stream())), N_SCORE, 8)
        one('ingest_f16', lambda: ingest(X16[:500_000], scaler), 500_000, D * 6)
        del sc_buf, fin
    del mass, jzw, Xs

    # ---- the other two training shapes of BASELINE.json at this N (weak scaling: per-GPU batch fixed):
    #   oe   = cfg-3 step: 5000 QCD jets paired with 5000 OoD (W) jets per GPU, OE_type 'KLD' (models.py:86-132), fused paired-tile kernel,
    #          peer-memory update at N > 1
    #   wide = cfg-5 step: FC_layers [1024, 512, 256, 64], batch 65536 per GPU (layered tcgen05 GEMM path); at N > 1 the 8.7 MB flat gradient
    #          is all-reduced by NCCL (the peer-memory update kernel needs one co-resident wave, which this net exceeds)
    extra = {}
    if not args.no_extra:
        n_ob = 40
        Xo16, _, _, _ = make_files(dev, 'W', n_ob * BATCH, 4321 + 7 * rank)
        Xo = ingest(Xo16, scaler); del Xo16                          # OoD sample (320 MB, larger than L2)
        voe = DeviceVAE(FC, D, 'KLD', args.precision, 'fixed', 0, local, BATCH)
        voe.set_params(init_params(FC, D, 0))
        p_oe = world > 1 and not os.environ.get('OEVAE_DP_NCCL') and voe.dp_connect()
        oe_ms = timed(make_step(voe, X, wts, BATCH, p_oe, grad, Xo, 1.0, 1.0), 5, 100)
        OE_ALG = 2 * (D * 4 + 4)                                  # 3,208 B per (bkg, OoD) pair
        extra['oe'] = {'metric': 'train pairs/s (cfg-3: bkg + OoD batch, OE_type KLD)', 'value': world * BATCH / (oe_ms * 1e-3), 'unit': 'pairs/s',
                       'ms_per_step': oe_ms, 'pairs_per_step_per_gpu': BATCH, 'update': 'peer-memory' if p_oe else ('nccl all-reduce' if world > 1 else 'single rank'),
                       'roofline': {'bound': 'hbm', 'achieved': OE_ALG * BATCH / (oe_ms * 1e-3) / 1e9, 'peak': hbm_peak, 'unit': 'GB/s',
                                    'frac': OE_ALG * BATCH / (oe_ms * 1e-3) / 1e9 / hbm_peak, 'alg_bytes_per_pair': OE_ALG}}
        del voe, Xo
        FCW, BW = [1024, 512, 256, 64], 65536
        vw = DeviceVAE(FCW, D, 'KLD', args.precision, 'fixed', 0, local, BW)
        vw.set_params(init_params(FCW, D, 0))
        p_w = world > 1 and not os.environ.get('OEVAE_DP_NCCL') and vw.dp_connect()
        gw = torch.empty(vw.n_params, dtype=torch.float32, device=dev)
        n_wb = n_jets // BW                                       # 15 distinct batches of the resident epoch
        w_ms = timed(make_step(vw, X[:n_wb * BW], torch.ones(BW, device=dev), BW, p_w, gw), 3, 10)
        hid = FCW[:-1]; L = FCW[-1]
        enc = D * hid[0] + sum(a_ * b_ for a_, b_ in zip(hid[:-1], hid[1:])) + 2 * hid[-1] * L
        dec = L * hid[-1] + sum(a_ * b_ for a_, b_ in zip(hid[::-1][:-1], hid[::-1][1:])) + hid[0] * D
        f_train = 3 * 2 * (enc + dec) - 2 * D * hid[0]            # 12,255,232 FLOP per jet
        ach = f_train * BW / (w_ms * 1e-3) / 1e12                 # per GPU
        extra['wide'] = {'metric': 'train jets/s (cfg-5: FC_layers [1024,512,256,64], batch 65536 per GPU)', 'value': world * BW / (w_ms * 1e-3), 'unit': UNIT,
                         'ms_per_step': w_ms, 'flop_per_jet': f_train, 'distinct_batches': n_wb,
                         'update': 'peer-memory' if p_w else ('nccl all-reduce of the %.1f MB flat gradient' % (vw.n_params * 4 / 1e6) if world > 1 else 'single rank'),
                         'roofline': {'bound': 'tensor', 'achieved': ach, 'peak': tc_peak, 'unit': 'TFLOP/s', 'frac': ach / tc_peak, 'peak_source': peak_src + ' (sustained cuBLAS bf16)',
                                      'note': 'algorithmic FLOP (one pass) per GPU; precision bf16x3 issues three bf16 MMAs per product'}}
        del vw, gw
        sums.zero_()

    # ---- end to end through the public API: models.train_model on a HOST sample.  The host holds what the reference's files hold
    # (float16 constituents, tools/merging.py:31) + weights; every step's batch is copied host->device from page-locked memory inside
    # the timed region and the running loss sums come back device->host after every step (written into page-locked memory by the update kernel).
    n_e2e = n_batches                                             # the whole 1M-jet epoch: 200 steps
    pin = lambda t: t.cpu().pin_memory().numpy()                  # page-locked host buffers, prepared before the timed region
    Xh = pin(X16[:n_e2e * BATCH]); wh = pin(torch.ones(n_e2e * BATCH))
    vh = pin(X16[:BATCH])
    bkg = {'constituents': Xh, 'weights': wh}
    vbk = {'constituents': vh, 'weights': wh[:BATCH]}
    model = VariationalAutoEncoder(FC, D, seed=0, precision=args.precision, device=local, max_batch=BATCH)
    e2e_reps, h2d_step = [], 0
    for rep in range(7):                                          # two repetitions warm the copy path; median of the other five
        st = {}
        sync()
        t0 = time.perf_counter()
        # N > 1: every rank feeds ITS OWN host shard (sharded_input): per-GPU batch 5000, gradients summed over ranks in the update kernel
        train_model(model, [(bkg, bkg)], [(vbk, vbk)], 'KLD', 1, BATCH, BETA, 0, 0, LR, verbose=False, scaler=scaler,
                    stats=st, sharded_input=True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        if rep >= 2: e2e_reps.append(max_over_ranks(dt))
        h2d_step = st.get('h2d_bytes', 0) / max(st.get('steps', 1), 1)
    e2e_ms = sorted(e2e_reps)[len(e2e_reps) // 2]
    e2e_val = world * n_e2e * BATCH / (e2e_ms * 1e-3)

    if rank != 0:
        return
    line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': ms / K,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': {'fp32': 'f32', 'bf16': 'bf16', 'bf16x3': 'bf16x3'}[args.precision],
            'data': 'synthetic', 'config': {'workload': workload_text(n_jets), 'global_batch': world * BATCH,
                                            'parallelism': 'dp%d' % world + (' (gradient exchange inside the update kernel over NVLink peer memory)' if peer else (' (NCCL all-reduce)' if world > 1 else '')),
                                            'l2': 'inputs larger than L2 (step i reads batch i %% %d: a fresh %.2f MB slice of the %.2f GB resident epoch)' % (n_batches, BATCH * D * 4 / 1e6, n_jets * D * 4 / 1e9),
                                            'precision': args.precision, 'mean_train_loss': loss_mean, 'numa': numa},
            'clocks': clk, 'gpu_launches': int(launches),
            'e2e': {'value': e2e_val, 'unit': UNIT, 'h2d_bytes_per_step': int(h2d_step), 'd2h_bytes_per_step': 32, 'ms_per_step': e2e_ms / n_e2e,
                    'reps_ms_per_step': [m_ / n_e2e for m_ in e2e_reps], 'steps': n_e2e,
                    'api': 'atlas_vae_b200.models.train_model(host float16 sample in page-locked memory, scaler=MaxAbs): 1 epoch of %d steps + its validation pass, wall clock, verbose=False; the running loss sums of every step are stored into page-locked host memory by its update kernel (32 B device->host per step), the host reads them once per load (nothing is printed)' % n_e2e},
            'roofline': roof, 'kernels': kernels[:8], 'scoring': scoring, 'strong': strong, 'small_kernels': small}
    if dp_parity is not None: line['dp_parity'] = dp_parity
    line.update(extra)
    if world == 1 and not args.no_cpu:
        jps, cores, dt, done = cpu_reference_steps(args.cpu_steps, 5, budget_s=20.0)
        line['cpu_baseline'] = {'value': jps, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                                'sample': '%d steps of %d jets (%.1f s): models.py:108-132,181-186 restated op-for-op in torch-CPU eager fp32' % (done, BATCH, dt)}
    os.write(real_stdout, (json.dumps(line) + '\n').encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--precision', default='bf16x3', choices=['fp32', 'bf16x3', 'bf16'])
    ap.add_argument('--cpu-steps', type=int, default=300)
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-extra', action='store_true', help='skip the cfg-3 (OE), cfg-5 (wide) and small-kernel legs')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
