import sys, os, numpy as np, torch
sys.path.insert(0, '.')
from atlas_vae_b200.engine import DeviceVAE
from oracle import vae_numpy as V, scoring as S
rng = np.random.default_rng(0)
FC, D, B = [80, 40, 10], 80, 512
flat = V.init_params(FC, D, 0)
X = rng.normal(0, 0.5, (B, D)).astype(np.float32); Xo = rng.normal(0.1, 0.7, (B, D)).astype(np.float32)
w = np.ones(B, np.float32); eps = rng.normal(size=(B, 10)).astype(np.float32)
def rel(a, b): return float(np.max(np.abs(a - b) / (np.abs(b) + 1e-12)))
dev = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'external', 0); dev.set_params(flat)
xhat0 = S.predict(flat, FC, D, X, [eps], np.float64); mae0 = np.mean(np.abs(X - xhat0), 1)
for tag, env in (('default', {}), ('static', {'OEVAE_SCORE_STATIC': '1'}), ('heads_t', {'OEVAE_SCORE_HEADS_T': '1'})):
    for k in ('OEVAE_SCORE_STATIC', 'OEVAE_SCORE_HEADS_T'): os.environ.pop(k, None)
    os.environ.update(env)
    o2 = dev.score(X, 1, eps, metrics=('MAE', 'Latent')); o5 = dev.score(X, 1, eps, metrics=('MSE', 'MAE', 'KLD', 'JSD', 'Latent'), want_xhat=True)
    print(tag, 'fresh params: MAE(2 metrics) rel', rel(o2['MAE'].cpu().numpy(), mae0), 'MAE(5 metrics) rel', rel(o5['MAE'].cpu().numpy(), mae0),
          'xhat max abs', float(np.abs(o5['xhat'].cpu().numpy() - xhat0).max()), flush=True)
for k in ('OEVAE_SCORE_STATIC', 'OEVAE_SCORE_HEADS_T'): os.environ.pop(k, None)
tr = V.Trainer(flat, FC, D, 'KLD')
for _ in range(2):
    pj = torch.empty((4, B), device='cuda')
    dev.train_step(X, w, None, None, eps, 1e-3, 1.0, 0.0, 0.0, pj)
    tr.step(X, w, None, None, eps, lr=1e-3, beta=1.0, lamb=0.0, margin=0.0)
th = dev.get_params()
print('params rel max', rel(th, tr.theta), 'abs max', float(np.abs(th - tr.theta).max()))
xh_dev = S.predict(th, FC, D, X, [eps], np.float64); xh_or = S.predict(tr.theta, FC, D, X, [eps], np.float64)
m_dev = np.mean(np.abs(X - xh_dev), 1); m_or = np.mean(np.abs(X - xh_or), 1)
o2 = dev.score(X, 1, eps, metrics=('MAE', 'Latent'))['MAE'].cpu().numpy()
print('after 2 steps: device MAE vs oracle(device params) rel', rel(o2, m_dev), ' vs oracle(oracle params) rel', rel(o2, m_or), ' oracle(dev params) vs oracle(oracle params)', rel(m_dev, m_or))
dev.set_params(th)
o3 = dev.score(X, 1, eps, metrics=('MAE', 'Latent'))['MAE'].cpu().numpy()
print('after set_params(get_params()): rel', rel(o3, m_dev))
dev2 = DeviceVAE(FC, D, 'KLD', 'bf16x3', 'external', 0); dev2.set_params(th)
o4 = dev2.score(X, 1, eps, metrics=('MAE', 'Latent'))['MAE'].cpu().numpy()
print('new engine with the trained params: rel', rel(o4, m_dev))
# training again, then eval_losses (fused forward) vs score
pj = torch.empty((4, B), device='cuda')
dev.train_step(X, w, None, None, eps, 1e-3, 1.0, 0.0, 0.0, pj)
th2 = dev.get_params(); m2 = np.mean(np.abs(X - S.predict(th2, FC, D, X, [eps], np.float64)), 1)
ev = dev.eval_losses(X, w, None, None, eps, 1.0, 0., 0.).cpu().numpy()[0]
o5 = dev.score(X, 1, eps, metrics=('MAE', 'Latent'))['MAE'].cpu().numpy()
print('after a 3rd step: eval_losses rec vs oracle rel', rel(ev, m2), ' score MAE rel', rel(o5, m2))
o6 = dev.score(X, 1, eps, metrics=('MAE', 'Latent'))['MAE'].cpu().numpy()
print('second score call rel', rel(o6, m2))
