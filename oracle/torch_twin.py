"""torch-CPU eager restatement of the reference training step (TEST INFRASTRUCTURE, see oracle/__init__.py).

Op-for-op twin of /root/reference/OE-VAE/models.py:7-132,181-186, written with torch autograd so that it is
an INDEPENDENT derivation of the gradient (the NumPy oracle's backward is hand-derived).  It deliberately
keeps the reference's inefficiencies -- the redundant second encoder(bkg) pass in get_OE_loss
(models.py:91), one clip + one Adam update per tensor (models.py:185-186) -- because it is also
the timed "reference CPU path" (bench.py cpu_baseline / --impl reference, kind="port"): TensorFlow itself is
not installable in this image.  PARITY UNPINNED (no TF to compare with).
"""
import numpy as np
import torch

from .vae_numpy import layer_dims


def clip_values(t, max_val=1e6):                       # models.py:243-246
    t = torch.where(torch.isfinite(t), t, torch.zeros_like(t))
    return torch.clamp(t, -max_val, max_val)


class TorchVAE:
    """Holds one tensor per Keras variable, in trainable_weights order."""

    def __init__(self, flat, FC_layers, input_dim, dtype=torch.float32):
        self.FC, self.D, self.dt = list(FC_layers), int(input_dim), dtype
        self.n_hidden = len(FC_layers) - 1
        self.vars, off = [], 0
        flat = np.asarray(flat)
        for i, o in layer_dims(FC_layers, input_dim):
            W = torch.tensor(flat[off:off + i * o].reshape(i, o), dtype=dtype, requires_grad=True); off += i * o
            b = torch.tensor(flat[off:off + o], dtype=dtype, requires_grad=True); off += o
            self.vars += [W, b]
        self.losses = []
        # Keras Adam slots (models.py:138)
        self.m = [torch.zeros_like(p) for p in self.vars]
        self.v = [torch.zeros_like(p) for p in self.vars]
        self.t = 0

    def flat(self):
        return np.concatenate([p.detach().numpy().ravel() for p in self.vars])

    def _dense(self, x, li, relu):
        y = x @ self.vars[2 * li] + self.vars[2 * li + 1]
        return torch.relu(y) if relu else y

    def encoder(self, x, eps):                          # models.py:30-36 + Sampling 9-15
        n = self.n_hidden
        for k in range(n):
            x = self._dense(x, k, True)
        mu = self._dense(x, n, False)
        lv = self._dense(x, n + 1, False)
        sigma = clip_values(torch.exp(lv / 2))
        z = eps * sigma + mu if eps is not None else None
        return mu, lv, z

    def decoder(self, z):                               # models.py:49-52
        n = self.n_hidden
        for k in range(n):
            z = self._dense(z, n + 2 + k, True)
        return self._dense(z, 2 * n + 2, False)

    def __call__(self, x, eps):                         # models.py:62-69
        mu, lv, z = self.encoder(x, eps)
        rec = clip_values(self.decoder(z))
        self.losses = [get_KLD_loss(mu, lv)]
        return rec


def get_reconstruction_loss(x, y, OE_type):             # models.py:72-78
    if OE_type in ('MSE', 'MSE-margin'):
        return torch.mean((x - y) ** 2, dim=-1)
    return torch.mean(torch.abs(x - y), dim=-1)


def get_KLD_loss(mu, lv):                               # models.py:81-85
    e = clip_values(torch.exp(lv))
    return -torch.mean(1 + lv - e - mu ** 2, dim=-1) / 2


def get_OE_loss(vae, X, X_OE, OE_type, margin, eps2, eps_ood):   # models.py:88-105
    if OE_type == 'KLD':
        mu, lv, _ = vae.encoder(X, eps2)                # redundant second pass kept on purpose (F11)
        mu_o, lv_o, _ = vae.encoder(X_OE, eps_ood)
        return torch.relu(get_KLD_loss(mu, lv) - get_KLD_loss(mu_o, lv_o) + margin)
    r = get_reconstruction_loss(X, vae(X, eps2), OE_type)
    r_o = get_reconstruction_loss(X_OE, vae(X_OE, eps_ood), OE_type)
    if OE_type in ('MSE', 'MAE'):
        return torch.sigmoid(r - r_o)
    return torch.relu(r - r_o + margin)


def get_losses(vae, X, w, X_OE, w_OE, eps, OE_type, beta, lamb, margin, eps2=None, eps_ood=None):   # models.py:108-132
    eps2 = eps if eps2 is None else eps2
    eps_ood = eps if eps_ood is None else eps_ood
    rec = vae(X, eps)
    loss_MSE = get_reconstruction_loss(X, rec, OE_type) * w
    loss_KLD = sum(vae.losses) * w * beta
    loss_OoD = get_OE_loss(vae, X, X_OE, OE_type, margin, eps2, eps_ood) * w_OE * lamb
    return loss_MSE, loss_KLD, loss_OoD, loss_MSE + loss_KLD + loss_OoD


def train_step(vae, X, w, X_OE, w_OE, eps, OE_type='KLD', beta=0., lamb=0., margin=0., lr=1e-3):
    """models.py:181-186: tape.gradient of the per-jet VECTOR == gradient of its sum; per-tensor clip; Keras Adam."""
    parts = get_losses(vae, X, w, X_OE, w_OE, eps, OE_type, beta, lamb, margin)
    grads = torch.autograd.grad(parts[3].sum(), vae.vars, allow_unused=True)
    grads = [torch.zeros_like(p) if g is None else clip_values(g) for g, p in zip(grads, vae.vars)]
    vae.t += 1
    t = vae.t
    alpha = lr * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
    with torch.no_grad():
        for p, g, m, v in zip(vae.vars, grads, vae.m, vae.v):      # one ResourceApplyAdam per variable
            m.mul_(0.9).add_(g, alpha=0.1)
            v.mul_(0.999).addcmul_(g, g, value=0.001)
            p.sub_(alpha * m / (v.sqrt() + 1e-7))
    return [q.detach() for q in parts], grads


def flat_grads(grads):
    return np.concatenate([g.detach().numpy().ravel() for g in grads])
