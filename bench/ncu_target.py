"""Small workload for ncu captures (one GPU): a few cfg-2 training steps, a cfg-3 (paired OE) step, one scoring pass with loss_mapping /
histogram, K0, fp16 ingest and the threshold search.  Sizes are kept small: ncu replays every captured kernel ~40 times.
  ncu --set full --clock-control none --import-source on -k regex:<kernel> -c <n> -o gpurun_out/<name> python bench/ncu_target.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from atlas_vae_b200 import _ffi, cuts as Cc, synthetic as S, utils as U
from atlas_vae_b200._ffi import ptr, check
from atlas_vae_b200.engine import DeviceVAE
from atlas_vae_b200.models import init_params

torch.cuda.set_device(0); dev = torch.device('cuda', 0)
FC, D, B = [80, 40, 20, 10], 400, 5000
st = lambda: torch.cuda.current_stream().cuda_stream
raw, _ = S.make_jets('QCD', 200_000, 200, seed=1, device=dev)
sc = None
x0, pt, m = U.preprocess_device(raw, 100, 4, None, True)                       # K0 (unscaled)
scaler = S.MaxAbs(S.maxabs_scale(x0[:100_000]))
X, pt, m = U.preprocess_device(raw, 100, 4, scaler, True)                      # K0 with the affine scaler
X16 = x0.to(torch.float16)
s_t = torch.as_tensor(scaler.scale_).to(dev)
Xi = torch.empty_like(X)
check(_ffi.lib().oevae_ingest_f16(ptr(X16), ptr(Xi), X16.shape[0], D, None, ptr(s_t), 0, st()))
w = torch.ones(X.shape[0], device=dev)
v = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'fixed', 0, 0, B); v.set_params(init_params(FC, D, 0))
sums = torch.zeros(4, dtype=torch.float64, device=dev)
for i in range(6):
    v.train_step(X[i * B:(i + 1) * B], w[:B], None, None, None, 1e-3, 1.0, 0., 0., None, sums)
for i in range(3):
    v.train_step(X[i * B:(i + 1) * B], w[:B], X[(i + 10) * B:(i + 11) * B], w[:B], None, 1e-3, 1.0, 1.0, 0.5, None, sums)
out = v.score(X, 1, None, metrics=('Latent', 'MAE', 'KLD', 'JSD'))
Cc.loss_mapping_device(out['MAE'], out['finite'])
jzw = torch.where(torch.rand(X.shape[0], device=dev) < 0.2, -1.0, 0.0)
edges = torch.as_tensor(U.get_idx(np.float32(m.max().item()), bin_size=2.5, min_val=0, integer=False, tuples=False)).to(dev)
counts = torch.zeros((2, edges.numel() - 1), dtype=torch.int64, device=dev); wsums = torch.zeros_like(counts, dtype=torch.float64)
check(_ffi.lib().oevae_hist(ptr(m), ptr(w), ptr(jzw), ptr(out['MAE']), ptr(out['finite']), X.shape[0], 0.3, ptr(edges), edges.numel(), 1.0, 0,
                            ptr(counts), ptr(wsums), 0, st()))
y = (jzw == 0).to(torch.uint8)
roc = Cc.roc_curve_device(y, out['MAE'], w)
print('cuts', roc.cuts([1e-2, 1e0, 1e1]))
torch.cuda.synchronize()
print('ncu target done: loss sums', sums.cpu().numpy())
