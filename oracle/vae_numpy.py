"""NumPy restatement of the OE-VAE training step (TEST INFRASTRUCTURE, see oracle/__init__.py).

Follows /root/reference/OE-VAE/models.py:
  Sampling.call            models.py:7-15     -> reparam()
  Encoder / Decoder        models.py:18-52    -> forward()
  VariationalAutoEncoder   models.py:55-69    -> forward() (clip of reconstruction, KLD add_loss)
  get_reconstruction_loss  models.py:72-78    -> rec_loss()
  get_KLD_loss             models.py:81-85    -> kld_loss()
  get_OE_loss              models.py:88-105   -> losses()
  get_losses               models.py:108-132  -> losses()
  train_model batch body   models.py:181-186  -> grad() / adam_step() / train_step()
  model_checkpoint         models.py:222-240  -> model_checkpoint()
  clip_values              models.py:243-246  -> clipv() / clip_mask()

The backward pass is hand-derived (the reference uses tf.GradientTape); it is the gradient of the
SUM over the batch of the per-jet total loss (models.py:132,183).  ``oracle/torch_twin.py`` derives
the same gradient by autograd; the two are cross-checked in tests/test_oracle.py.

PARITY UNPINNED for this file: TensorFlow 2.1 is not installable here, the reference has no golden
vectors.  Everything is computed in ``dtype`` (float64 "truth" or float32
"reference dtype"); the noise eps is an explicit input.
"""
import numpy as np

OE_TYPES = ('KLD', 'MSE', 'MAE', 'MSE-margin', 'MAE-margin')
CLIP = 1e6


# ----------------------------------------------------------------------------- layout
def layer_dims(FC_layers, input_dim):
    """(in,out) of every Dense in Keras trainable_weights order (models.py:22-28,43-48)."""
    hidden, L = list(FC_layers[:-1]), int(FC_layers[-1])
    dims, d = [], int(input_dim)
    for h in hidden:                       # encoder hidden
        dims.append((d, h)); d = h
    dims.append((d, L))                    # dense_mean
    dims.append((d, L))                    # dense_log_var
    d = L
    for h in hidden[::-1]:                 # decoder hidden (mirrored)
        dims.append((d, h)); d = h
    dims.append((d, int(input_dim)))       # dense_output
    return dims


def param_layout(FC_layers, input_dim):
    """[(name, shape, offset)] of the flat fp32 parameter vector (kernel [in,out] row-major, then bias)."""
    out, off = [], 0
    n = len(FC_layers) - 1
    names = ['enc%d' % i for i in range(n)] + ['mean', 'log_var'] + ['dec%d' % i for i in range(n)] + ['out']
    for name, (i, o) in zip(names, layer_dims(FC_layers, input_dim)):
        out.append((name + '.kernel', (i, o), off)); off += i * o
        out.append((name + '.bias', (o,), off)); off += o
    return out, off


def n_params(FC_layers, input_dim):
    return param_layout(FC_layers, input_dim)[1]


def unflatten(flat, FC_layers, input_dim):
    lay, n = param_layout(FC_layers, input_dim)
    assert flat.shape == (n,)
    ts = [flat[o:o + int(np.prod(s))].reshape(s) for _, s, o in lay]
    return [(ts[2 * i], ts[2 * i + 1]) for i in range(len(ts) // 2)]


def init_params(FC_layers, input_dim, seed=0):
    """Initial flat fp32 parameters with the reference's initialisers (models.py:23-28,44-48):
    hidden kernels he_normal (truncated normal, std=sqrt(2/fan_in)/0.8796...), hidden biases N(0,1),
    heads / output: glorot-uniform kernels, zero biases.  The stream is NumPy Philox, not TF's."""
    rng = np.random.Generator(np.random.Philox(seed))
    n = len(FC_layers) - 1
    dims = layer_dims(FC_layers, input_dim)
    hidden_idx = set(range(n)) | set(range(n + 2, 2 * n + 2))
    parts = []
    for li, (i, o) in enumerate(dims):
        if li in hidden_idx:
            std = np.sqrt(2.0 / i) / 0.87962566103423978
            k = rng.standard_normal((i, o))
            bad = np.abs(k) > 2
            while bad.any():                       # truncated at +-2 sigma by resampling
                k[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(k) > 2
            k *= std
            b = rng.standard_normal(o)
        else:
            lim = np.sqrt(6.0 / (i + o))
            k = rng.uniform(-lim, lim, (i, o))
            b = np.zeros(o)
        parts += [k.ravel(), b]
    return np.concatenate(parts).astype(np.float32)


# ----------------------------------------------------------------------------- primitives
def clipv(t, M=CLIP):
    """clip_values (models.py:243-246): non-finite -> 0, then clamp to [-M, M]."""
    t = np.where(np.isfinite(t), t, 0)
    return np.clip(t, -M, M)


def clip_mask(t, M=CLIP):
    """Gradient mask of clip_values: tf.where routes the gradient to the finite branch and
    clip_by_value passes it on [-M, M] (inclusive)  (SURVEY appendix A/B)."""
    with np.errstate(invalid='ignore'):
        return (np.isfinite(t) & (np.abs(t) <= M)).astype(t.dtype)


def kld_loss(mu, lv):
    """get_KLD_loss (models.py:81-85): MEAN over latent dims, /2."""
    with np.errstate(over='ignore'):
        e = clipv(np.exp(lv))
    return -np.mean(1 + lv - e - np.square(mu), axis=-1) / 2


def rec_loss(x, y, OE_type):
    """get_reconstruction_loss (models.py:72-78): mean over features."""
    if OE_type in ('MSE', 'MSE-margin'):
        return np.mean(np.square(x - y), axis=-1)
    return np.mean(np.abs(x - y), axis=-1)


# ----------------------------------------------------------------------------- forward
def forward(layers, x, eps, n_hidden, encoder_only=False):
    """Full VAE forward; returns a cache with every tensor needed by backward()."""
    dt = x.dtype
    c = {'a': [x]}
    a = x
    for k in range(n_hidden):
        W, b = layers[k]
        a = np.maximum(a @ W + b, 0)
        c['a'].append(a)
    Wm, bm = layers[n_hidden]
    Wv, bv = layers[n_hidden + 1]
    mu = a @ Wm + bm
    lv = a @ Wv + bv
    c['mu'], c['lv'] = mu, lv
    c['kld'] = kld_loss(mu, lv)
    if encoder_only:
        return c
    with np.errstate(over='ignore'):
        s_raw = np.exp(lv / 2)
    s = clipv(s_raw)
    z = mu + s * eps.astype(dt)                      # tf.random.normal(mean=mu, stddev=sigma): eps*sigma+mu
    c['s_raw'], c['eps'], c['z'] = s_raw, eps.astype(dt), z
    d = z
    c['d'] = [z]
    for k in range(n_hidden):
        V, cb = layers[n_hidden + 2 + k]
        d = np.maximum(d @ V + cb, 0)
        c['d'].append(d)
    Vo, co = layers[2 * n_hidden + 2]
    y_raw = d @ Vo + co
    c['y_raw'] = y_raw
    c['y'] = clipv(y_raw)
    return c


def backward(layers, c, g_rec, g_kld, OE_type, n_hidden, grads):
    """Accumulate into ``grads`` (list of [gW, gb]) the gradient of  sum_i g_rec_i*rec_i + g_kld_i*kld_i
    for one pass whose cache is ``c``.  g_rec None => encoder-only pass (no decoder / z path)."""
    mu, lv = c['mu'], c['lv']
    dt = mu.dtype
    L = mu.shape[1]
    if g_rec is not None:
        x, y, y_raw = c['a'][0], c['y'], c['y_raw']
        D = x.shape[1]
        if OE_type in ('MSE', 'MSE-margin'):
            g_y = g_rec[:, None] * 2 * (y - x) / D
        else:
            g_y = g_rec[:, None] * np.sign(y - x) / D
        g = g_y * clip_mask(y_raw)
        for k in range(n_hidden, -1, -1):            # output layer, then decoder hidden in reverse
            li = n_hidden + 2 + k
            inp = c['d'][k]
            grads[li][0] += inp.T @ g
            grads[li][1] += g.sum(0)
            g = g @ layers[li][0].T
            if k > 0:
                g = g * (inp > 0)
        g_z = g
    else:
        g_z = None
    with np.errstate(over='ignore', invalid='ignore'):
        e_raw = np.exp(lv)
        g_mu = g_kld[:, None] * mu / L
        g_lv = g_kld[:, None] * (-(1 - np.where(clip_mask(e_raw) > 0, e_raw, 0)) / (2 * L))
        if g_z is not None:
            s_raw = c['s_raw']
            g_mu = g_mu + g_z
            g_lv = g_lv + g_z * c['eps'] * np.where(clip_mask(s_raw) > 0, s_raw / 2, 0)
    a_n = c['a'][n_hidden]
    grads[n_hidden][0] += a_n.T @ g_mu
    grads[n_hidden][1] += g_mu.sum(0)
    grads[n_hidden + 1][0] += a_n.T @ g_lv
    grads[n_hidden + 1][1] += g_lv.sum(0)
    g = (g_mu @ layers[n_hidden][0].T + g_lv @ layers[n_hidden + 1][0].T)
    for k in range(n_hidden - 1, -1, -1):
        g = g * (c['a'][k + 1] > 0)
        grads[k][0] += c['a'][k].T @ g
        grads[k][1] += g.sum(0)
        if k > 0:
            g = g @ layers[k][0].T
    return grads


def _sigmoid(t):
    return 1 / (1 + np.exp(-t))


def losses_and_grad(flat, FC_layers, input_dim, x, w, x_ood=None, w_ood=None, eps=None, eps2=None,
                    eps_ood=None, OE_type='KLD', beta=0., lamb=0., margin=0., dtype=np.float64,
                    want_grad=True):
    """get_losses (models.py:108-132) + tape.gradient of the SUM (models.py:183), un-clipped.

    eps  : noise of the vae(bkg) pass  (models.py:122)
    eps2 : noise of the second vae(bkg) pass inside get_OE_loss for the non-KLD types (models.py:98);
           defaults to eps (reference seed=0 behaviour)
    eps_ood : noise of vae(OoD) (models.py:99); defaults to eps.
    Returns ((rec_w, kld_w, oe_w, total) each [B], flat gradient or None).
    """
    assert OE_type in OE_TYPES
    dt = dtype
    n_hidden = len(FC_layers) - 1
    layers = [(W.astype(dt), b.astype(dt)) for W, b in unflatten(np.asarray(flat), FC_layers, input_dim)]
    x = np.asarray(x, dt); w = np.asarray(w, dt)
    B = x.shape[0]
    L = FC_layers[-1]
    eps = np.zeros((B, L), dt) if eps is None else np.asarray(eps, dt)
    have_ood = x_ood is not None
    if have_ood:
        x_ood = np.asarray(x_ood, dt); w_ood = np.asarray(w_ood, dt)
    c1 = forward(layers, x, eps, n_hidden)
    rec = rec_loss(x, c1['y'], OE_type)
    kld = c1['kld']
    loss_rec = rec * w
    loss_kld = kld * w * dt(beta)
    g_rec1 = w.copy()
    g_kld1 = w * dt(beta)
    passes = []
    if have_ood:
        if OE_type == 'KLD':
            co = forward(layers, x_ood, None, n_hidden, encoder_only=True)
            delta = kld - co['kld'] + dt(margin)
            oe = np.maximum(delta, 0)
            act = (delta > 0).astype(dt)
            coef = dt(lamb) * w_ood * act
            g_kld1 = g_kld1 + coef                     # second encoder(bkg) pass == first (deterministic, F11)
            passes.append((co, None, -coef))
        else:
            eps2 = eps if eps2 is None else np.asarray(eps2, dt)
            eps_ood = eps if eps_ood is None else np.asarray(eps_ood, dt)
            c2 = forward(layers, x, eps2, n_hidden)
            co = forward(layers, x_ood, eps_ood, n_hidden)
            r2 = rec_loss(x, c2['y'], OE_type)
            ro = rec_loss(x_ood, co['y'], OE_type)
            if OE_type in ('MSE', 'MAE'):
                oe = _sigmoid(r2 - ro)
                dphi = oe * (1 - oe)
            else:
                delta = r2 - ro + dt(margin)
                oe = np.maximum(delta, 0)
                dphi = (delta > 0).astype(dt)
            coef = dt(lamb) * w_ood * dphi
            zero = np.zeros(B, dt)
            passes.append((c2, coef, zero))
            passes.append((co, -coef, zero))
        loss_oe = oe * w_ood * dt(lamb)
    else:
        loss_oe = np.zeros(B, dt)
    total = loss_rec + loss_kld + loss_oe
    out = (loss_rec, loss_kld, loss_oe, total)
    if not want_grad:
        return out, None
    grads = [[np.zeros_like(W), np.zeros_like(b)] for W, b in layers]
    backward(layers, c1, g_rec1, g_kld1, OE_type, n_hidden, grads)
    for c, gr, gk in passes:
        backward(layers, c, gr, gk, OE_type, n_hidden, grads)
    gflat = np.concatenate([np.concatenate([gW.ravel(), gb]) for gW, gb in grads])
    return out, gflat


# ----------------------------------------------------------------------------- optimiser
def adam_step(theta, m, v, g, t, lr, dtype=np.float64, b1=0.9, b2=0.999, eps=1e-7):
    """Keras TF-2.1 Adam (models.py:138,186; hyper-parameters spelled out at jet-ID/models.py:87) applied
    to gradients sanitised with clip_values (models.py:185).  epsilon sits OUTSIDE the bias correction.
    t is the 1-based global step."""
    dt = dtype
    g = clipv(np.asarray(g, dt))
    m = dt(b1) * m + dt(1 - b1) * g
    v = dt(b2) * v + dt(1 - b2) * g * g
    alpha = dt(lr) * np.sqrt(dt(1) - dt(b2) ** t) / (dt(1) - dt(b1) ** t)
    theta = theta - alpha * m / (np.sqrt(v) + dt(eps))
    return theta, m, v


class Trainer:
    """State-carrying restatement of the batch loop body (models.py:177-190)."""

    def __init__(self, flat, FC_layers, input_dim, OE_type='KLD', dtype=np.float64):
        self.FC, self.D, self.OE_type, self.dt = list(FC_layers), input_dim, OE_type, dtype
        self.theta = np.asarray(flat, dtype).copy()
        self.m = np.zeros_like(self.theta); self.v = np.zeros_like(self.theta); self.t = 0

    def step(self, x, w, x_ood=None, w_ood=None, eps=None, lr=1e-3, beta=0., lamb=0., margin=0., **kw):
        out, g = losses_and_grad(self.theta, self.FC, self.D, x, w, x_ood, w_ood, eps, OE_type=self.OE_type,
                                 beta=beta, lamb=lamb, margin=margin, dtype=self.dt, **kw)
        self.t += 1
        self.theta, self.m, self.v = adam_step(self.theta, self.m, self.v, g, self.t, lr, self.dt)
        self.last_grad = g
        return out


# ----------------------------------------------------------------------------- epoch-level host logic
def model_checkpoint(history, lr, count, metric='Train loss', patience=3, factor=2, min_delta=1e-3, min_lr=1e-4):
    """models.py:222-240 as a pure function: returns (save_weights?, new_lr, new_count or None=stop)."""
    h = history[metric]
    save = False
    if h[-1] < np.min(h[:-1]) - min_delta:
        save, count = True, 0
    elif h[-1] > np.min(h[-(patience + 1):-1]) - min_delta:
        count += 1
    if count >= patience:
        if lr < min_lr:
            count = None
        else:
            lr = lr / factor
            count = 0
    return save, lr, count
