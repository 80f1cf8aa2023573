import sys, os, subprocess, threading, time, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params
FC, D = [80, 40, 20, 10], 400
N = 148 * 64 * 128
X = torch.rand(N, D, device='cuda') * 0.3
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, None, 5000); dev.set_params(init_params(FC, D, 0))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
M4 = ('MAE', 'KLD', 'JSD', 'Latent')
def t(n, metrics, reps=5):
    best = 1e9
    for r in range(reps):
        e0.record(); dev.score(X[:n], metrics=metrics, want_finite=False); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
for flag in ('0', '1'):
    os.environ['OEVAE_SCORE_HEADS_T'] = flag
    prev = None
    for k in (4, 8, 16, 32, 64, 128):
        ms = t(148 * 64 * k, M4)
        inc = '' if prev is None else ' incremental cycles/tile @1965MHz %.0f' % ((ms - prev[1]) * 1e-3 * 1.965e9 / (k - prev[0]))
        print('heads_t', flag, 'tiles/CTA', k, 'ms', round(ms, 4), inc, flush=True)
        prev = (k, ms)
# clocks under a sustained scoring loop
samples = []
stop = False
def poll():
    while not stop:
        try:
            o = subprocess.run(['nvidia-smi', '--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active', '--format=csv,noheader'], capture_output=True, text=True, timeout=5).stdout.strip()
            samples.append(o)
        except Exception as ex:
            samples.append(str(ex))
        time.sleep(0.2)
th = threading.Thread(target=poll); th.start()
t0 = time.time()
while time.time() - t0 < 4.0:
    dev.score(X, metrics=M4, want_finite=False)
torch.cuda.synchronize()
stop = True; th.join()
print('clock samples under load:', samples[::3])
