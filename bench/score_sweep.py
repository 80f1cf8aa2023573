"""Scoring pass (oevae_score, fused SCORE kernel) timed against the number of resident jets: jets/s and SM-cycle cost per 64-jet tile.
    python bench/score_sweep.py            (OEVAE_SCORE_STATIC=1 for the round-robin tile order)"""
import os, sys, subprocess, threading, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params

FC, D = [80, 40, 20, 10], 400
dev = torch.device('cuda', 0)
v = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, 0, 5000); v.set_params(init_params(FC, D, 0))
g = torch.Generator(device=dev); g.manual_seed(0)
_pre = [torch.empty(int(float(os.environ['SWEEP_PREALLOC']) * (1 << 28)), device=dev) for _ in range(1)] if os.environ.get('SWEEP_PREALLOC') else None   # GiB of other live allocations
NMAX = 5_000_000
if os.environ.get('SWEEP_DATA', 'randn') == 'jets':        # bench.py's scoring sample: synthetic QCD + top jets, MaxAbs-scaled float16 rows
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench as Bn
    from atlas_vae_b200 import synthetic as S
    NMAX = 2_500_000
    Xq16, _, _, _ = Bn.make_files(dev, 'QCD', NMAX - 250_000, 555)
    Xt16, _, _, _ = Bn.make_files(dev, 'top', 250_000, 777)
    scaler = S.MaxAbs(S.maxabs_scale(Xq16[:100_000].float()))
    X = torch.empty((NMAX, D), device=dev)
    X[:NMAX - 250_000] = Bn.ingest(Xq16, scaler); X[NMAX - 250_000:] = Bn.ingest(Xt16, scaler)
    del Xq16, Xt16
    print('fraction of exact zeros in X: %.3f' % float((X[:100000] == 0).float().mean()))
else:
    X = torch.empty((NMAX, D), device=dev)
    for lo in range(0, NMAX, 500_000): X[lo:lo + 500_000] = torch.randn((500_000, D), device=dev, generator=g) * 0.3
metrics = ('MAE', 'KLD', 'JSD', 'Latent')
clk = []
def poll():
    while not stop:
        try: clk.append(int(subprocess.run(['nvidia-smi', '--query-gpu=clocks.sm', '--format=csv,noheader,nounits', '-i', '0'], capture_output=True, text=True).stdout.split()[0]))
        except Exception: pass
        time.sleep(0.05)
for n in (250_000, 500_000, 1_000_000, 1_500_000, 2_500_000, 5_000_000):
    if n > NMAX: continue
    x = X[:n]
    for _ in range(2): v.score(x, metrics=metrics)
    torch.cuda.synchronize()
    stop = False; clk.clear(); th = threading.Thread(target=poll); th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(3, 10_000_000 // n)
    e0.record()
    for _ in range(reps): v.score(x, metrics=metrics)
    e1.record(); torch.cuda.synchronize()
    stop = True; th.join()
    ms = e0.elapsed_time(e1) / reps
    tiles_per_sm = (n / 64) / 148
    print('%9d jets: %7.3f ms  %6.1f M jets/s  %5.1f k SM-cycles per tile at 1.83 GHz  (sm clock samples: %s)'
          % (n, ms, n / ms / 1e3, ms * 1e-3 * 1.83e9 / tiles_per_sm / 1e3, sorted(set(clk))[:4]), flush=True)

if os.environ.get('SWEEP_TRAIN'):                                # the same pass with a trained model (bench.py scores after its training legs)
    w = torch.ones(5000, device=dev)
    for i in range(int(os.environ['SWEEP_TRAIN'])): v.train_step(X[(i % 200) * 5000:(i % 200 + 1) * 5000], w, None, None, None, 1e-3, 1.0, 0., 0.)
    n = min(NMAX, 2_500_000); x = X[:n]
    for _ in range(2): v.score(x, metrics=metrics)
    torch.cuda.synchronize(); e0.record()
    for _ in range(4): v.score(x, metrics=metrics)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 4
    print('after %s training steps: %9d jets: %7.3f ms  %6.1f M jets/s' % (os.environ['SWEEP_TRAIN'], n, ms, n / ms / 1e3))
