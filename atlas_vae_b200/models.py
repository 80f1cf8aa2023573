"""Drop-in for /root/reference/OE-VAE/models.py: same names, argument meaning and side effects, with every
numerical operation executed by liboevae.so (hand-written sm_100a CUDA) instead of TensorFlow-Keras.

  VariationalAutoEncoder   models.py:55-69   (+ Encoder 18-36, Decoder 39-52, Sampling 7-15)
  get_losses               models.py:108-132
  train_model              models.py:135-219
  model_checkpoint         models.py:222-240

Differences that are deliberate: ``train_model`` saves the weights after a single-epoch run too
(the reference only saves for epoch>0 yet vae.py:141 reloads ``model_out``), and weights are stored as an
.npz archive with Keras-style variable names (h5py is not available here).
"""
import os
import pickle
import sys
import time

import numpy as np
import torch

from .engine import DeviceVAE
from . import _ffi
from . import dist as _dist


class HostArray(np.ndarray):
    """ndarray with the ``.numpy()`` accessor reference callers use on TF tensors."""
    def numpy(self):
        return np.asarray(self)


def _host(t):
    return t.detach().cpu().numpy().view(HostArray)


def init_params(FC_layers, input_dim, seed=0):
    """Reference initialisers (models.py:23-28,44-48): hidden Dense = he_normal kernel (truncated normal,
    stddev sqrt(2/fan_in)/0.8796...) + N(0,1) bias; dense_mean / dense_log_var / dense_output = Keras
    defaults (glorot-uniform kernel, zero bias).  The random stream is NumPy's Philox, not TensorFlow's."""
    rng = np.random.Generator(np.random.Philox(seed))
    hidden, L = list(FC_layers[:-1]), int(FC_layers[-1])
    n = len(hidden)
    dims, d = [], int(input_dim)
    for h in hidden:
        dims.append((d, h, True)); d = h
    dims += [(d, L, False), (d, L, False)]
    d = L
    for h in hidden[::-1]:
        dims.append((d, h, True)); d = h
    dims.append((d, int(input_dim), False))
    parts = []
    for fan_in, fan_out, is_hidden in dims:
        if is_hidden:
            k = rng.standard_normal((fan_in, fan_out))
            bad = np.abs(k) > 2
            while bad.any():
                k[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(k) > 2
            k *= np.sqrt(2.0 / fan_in) / 0.87962566103423978
            b = rng.standard_normal(fan_out)
        else:
            lim = np.sqrt(6.0 / (fan_in + fan_out))
            k = rng.uniform(-lim, lim, (fan_in, fan_out))
            b = np.zeros(fan_out)
        parts += [k.ravel(), b]
    return np.concatenate(parts).astype(np.float32)


def _weight_names(n_hidden):
    names = []
    for i in range(n_hidden):
        names.append('encoder/dense%s' % ('' if i == 0 else '_%d' % i))
    names += ['encoder/dense_mean', 'encoder/dense_log_var']
    for i in range(n_hidden):
        names.append('decoder/dense_%d' % (n_hidden + i))
    names.append('decoder/dense_output')
    return names


class Adam:
    """Stand-in for tf.keras.optimizers.Adam(learning_rate=lr) (models.py:138): the moments live in the device
    handle; this object only carries the mutable learning rate that model_checkpoint halves (models.py:237)."""
    def __init__(self, learning_rate=1e-3):
        self.learning_rate = float(learning_rate)


class _Encoder:
    def __init__(self, vae):
        self._vae = vae

    def __call__(self, x, eps=None):
        """(z_mean, z_log_var, z)  -- models.py:30-36"""
        mu, lv, z = self._vae.dev.encode(x, eps)
        return _host(mu), _host(lv), (None if z is None else _host(z))


class VariationalAutoEncoder:
    """models.py:55-69.  ``seed=0`` reproduces the reference's re-seeded sampling (the same eps tensor on every
    forward call of a given shape); ``seed=None`` draws fresh Philox noise per call.
    Extra keyword-only knobs (not in the reference): precision, noise ('external' for explicit eps), device,
    init_seed (weight initialisation), max_batch (workspace hint)."""

    def __init__(self, FC_layers, input_dim, activation='relu', seed=None, name='autoencoder', *,
                 precision='fp32', noise=None, device=None, init_seed=0, max_batch=0):
        if activation != 'relu':
            raise ValueError("only activation='relu' (the reference default, models.py:57) is implemented")
        self.name = name
        self.FC_layers = [int(v) for v in FC_layers]
        self.input_dim = int(input_dim)
        self.seed = seed
        if noise is None:
            noise = 'per_step' if seed is None else 'fixed'
        key = int.from_bytes(os.urandom(8), 'little') if seed is None else int(seed)
        self.dev = DeviceVAE(self.FC_layers, self.input_dim, 'KLD', precision, noise, key, device, max_batch)
        self.dev.set_params(init_params(self.FC_layers, self.input_dim, init_seed))
        self.encoder = _Encoder(self)
        self.losses = []

    # ---- Keras-like surface
    def __call__(self, inputs, eps=None):
        """Reconstruction (clipped, models.py:66); sets ``.losses`` to [per-jet KLD] (models.py:68)."""
        out = self.dev.score(inputs, 1, eps, metrics=('Latent',), want_xhat=True, want_finite=False)
        # add_loss holds the raw KLD; 'Latent' is the same quantity with non-finite values zeroed (utils.py:384)
        self.losses = [_host(out['Latent'])]
        return _host(out['xhat'])

    def predict(self, X, batch_size=int(1e4), verbose=0, eps=None):
        """model.predict (vae.py:165).  batch_size is accepted for compatibility; chunking is internal."""
        return self.__call__(X, eps)

    @property
    def trainable_weights(self):
        flat = self.dev.get_params()
        out = []
        for w_off, b_off, i, o in self.dev.layout():
            out.append(flat[w_off:w_off + i * o].reshape(i, o).view(HostArray))
            out.append(flat[b_off:b_off + o].view(HostArray))
        return out

    def get_weights(self):
        return [np.asarray(w) for w in self.trainable_weights]

    def set_weights(self, weights):
        self.dev.set_params(np.concatenate([np.asarray(w, np.float32).ravel() for w in weights]))

    def save_weights(self, path):
        """models.py:227.  .npz archive keyed by Keras-style variable names, plus the flat vector."""
        ws = self.get_weights()
        names = _weight_names(len(self.FC_layers) - 1)
        d = {}
        for i, nme in enumerate(names):
            d[nme + '/kernel:0'] = ws[2 * i]; d[nme + '/bias:0'] = ws[2 * i + 1]
        d['FC_layers'] = np.array(self.FC_layers); d['input_dim'] = np.array(self.input_dim)
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with open(path, 'wb') as f:
            np.savez(f, **d)

    def load_weights(self, path):
        """vae.py:101,141"""
        with np.load(path) as d:
            if [int(v) for v in d['FC_layers']] != self.FC_layers or int(d['input_dim']) != self.input_dim:
                raise ValueError('checkpoint %s was written for FC_layers=%s input_dim=%d' % (path, list(d['FC_layers']), int(d['input_dim'])))
            names = _weight_names(len(self.FC_layers) - 1)
            ws = []
            for nme in names:
                ws += [d[nme + '/kernel:0'], d[nme + '/bias:0']]
        self.set_weights(ws)


# ------------------------------------------------------------------------------------------------ get_losses
def _pick_X(sample):
    """models.py:110-118"""
    if 'constituents' in sample and 'HLVs' in sample:
        return np.hstack([sample['constituents'], sample['HLVs']])
    if 'constituents' in sample:
        return sample['constituents']
    if 'HLVs' in sample:
        return sample['HLVs']
    raise KeyError("sample has neither 'constituents' nor 'HLVs'")


def get_losses(vae, data, OE_type, beta, lamb, margin, eps=None):
    """models.py:108-132 -> (loss_MSE, loss_KLD, loss_OoD, total), each [B]."""
    bkg, ood = data
    vae.dev.set_oe_type(OE_type)
    pj = vae.dev.eval_losses(_pick_X(bkg), bkg['weights'], _pick_X(ood), ood['weights'], eps,
                             float(beta), float(lamb), float(margin))
    pj = _host(pj)
    return pj[0], pj[1], pj[2], pj[3]


# ------------------------------------------------------------------------------------------------ staging
class _BatchStager:
    """Host (NumPy) load -> device batches.  The arrays of a load are page-locked in place once
    (cudaHostRegister) and every batch is copied host->device on a side stream, two batches ahead of the
    compute stream, so the training loop never waits on PCIe unless PCIe is the bottleneck."""

    def __init__(self, arrays, batch_size, device, shard=(0, 1)):
        self.device = device
        self.arrays = [np.ascontiguousarray(a, np.float32) for a in arrays]
        self.n = len(self.arrays[0])
        self.batch_size = int(batch_size)
        self.n_batches = int(np.ceil(self.n / self.batch_size))
        self.rank, self.world = shard
        self.copy_stream = torch.cuda.Stream(device=device)
        self._registered = []
        for a in self.arrays:
            if a.nbytes and _ffi.lib().oevae_host_register(a.ctypes.data, a.nbytes) == 0:
                self._registered.append(a)
        self.h2d_bytes = 0

    def bounds(self, b):
        lo, hi = b * self.batch_size, min((b + 1) * self.batch_size, self.n)
        if self.world > 1:                      # contiguous shard of the global batch
            n = hi - lo
            s0, s1 = lo + (n * self.rank) // self.world, lo + (n * (self.rank + 1)) // self.world
            return s0, s1, lo
        return lo, hi, lo

    def _issue(self, b):
        lo, hi, _ = self.bounds(b)
        out = []
        with torch.cuda.stream(self.copy_stream):
            for a in self.arrays:
                src = torch.from_numpy(a[lo:hi])
                dst = torch.empty(src.shape, dtype=torch.float32, device=self.device)
                dst.copy_(src, non_blocking=True)
                self.h2d_bytes += src.numel() * 4
                out.append(dst)
            ev = torch.cuda.Event(); ev.record(self.copy_stream)
        return out, ev

    def __iter__(self):
        pending = [self._issue(b) for b in range(min(2, self.n_batches))]
        for b in range(self.n_batches):
            tensors, ev = pending.pop(0)
            if b + 2 < self.n_batches:
                pending.append(self._issue(b + 2))
            torch.cuda.current_stream().wait_event(ev)
            for t in tensors:
                t.record_stream(torch.cuda.current_stream())
            yield b, tensors

    def close(self):
        torch.cuda.synchronize(self.device)
        for a in self._registered:
            _ffi.lib().oevae_host_unregister(a.ctypes.data)
        self._registered = []


# ------------------------------------------------------------------------------------------------ train_model
def train_model(vae, train_sample, valid_sample, OE_type='KLD', n_epochs=1, batch_size=5000,
                beta=0, lamb=0, margin=0, lr=1e-3, hist_file=None, model_in=None, model_out=None,
                patience=3, verbose=True):
    """models.py:135-219.  Same control flow, history schema, print cadence and LR-patience logic; the batch
    body (get_losses + tape.gradient + clip + Adam + metric updates, models.py:180-190) is one C-ABI call.
    Under torchrun (WORLD_SIZE>1) every global batch is split contiguously across ranks and the flat
    gradient is all-reduced with SUM before the update."""
    dev = vae.dev
    dev.set_oe_type(OE_type)
    rank, world = _dist.rank_world()
    out = sys.stdout if (verbose and rank == 0) else open(os.devnull, 'w')
    optimizer = Adam(learning_rate=lr)
    train_len, valid_len = len(train_sample), len(valid_sample)
    if train_len == 1:
        try: train_sample = [train_sample[0]]
        except Exception: pass
    if valid_len == 1:
        try: valid_sample = [valid_sample[0]]
        except Exception: pass
    print('STARTING TRAINING (generator ' + ('OFF' if train_len == 1 else 'ON') + ')', file=out)
    history = {'MSE': []}
    if beta != 0: history['KLD'] = []
    if lamb != 0: history['OE'] = []
    history = {**history, 'Train loss': [], 'Valid loss': []}
    if hist_file is not None and os.path.isfile(hist_file) and model_in and os.path.isfile(model_in):
        history = pickle.load(open(hist_file, 'rb'))
    sums = torch.zeros(4, dtype=torch.float64, device=dev.device)           # keras.metrics.Mean numerators
    vsum = torch.zeros(4, dtype=torch.float64, device=dev.device)
    grad = torch.empty(dev.n_params, dtype=torch.float32, device=dev.device) if world > 1 else None
    # n_gpus > 1: the update kernel exchanges the per-rank gradient sums over NVLink peer memory (one node, NCCL backend);
    # otherwise (gloo test boxes, nets too large for a co-resident launch) SUM all-reduce of the flat gradient
    peer = world > 1 and torch.distributed.get_backend() == 'nccl' and (getattr(dev, 'dp', False) or dev.dp_connect())
    total_batches = 0; count = 0
    beta, lamb, margin = float(beta), float(lamb), float(margin)
    for epoch in range(n_epochs):
        sum_batches = 0; n_seen = 0; start_time = time.time()
        print('\nEpoch %d/%d:' % (epoch + 1, n_epochs), file=out)
        sums.zero_(); vsum.zero_()
        losses = {}
        for train_idx in range(len(train_sample)):
            bkg, ood = train_sample[train_idx]
            stager = _BatchStager([_pick_X(bkg), bkg['weights'], _pick_X(ood), ood['weights']], batch_size, dev.device, (rank, world))
            if epoch == 0: total_batches += stager.n_batches
            try:
                for batch_idx, (x, w, xo, wo) in stager:
                    sum_batches += 1
                    lo, hi, glo = stager.bounds(batch_idx)
                    if world == 1:
                        dev.train_step(x, w, xo, wo, None, optimizer.learning_rate, beta, lamb, margin, None, sums)
                    elif peer:
                        dev.set_noise_offset(lo - glo)
                        dev.train_step_dp(x, w, xo, wo, None, optimizer.learning_rate, beta, lamb, margin, None, sums)
                    else:
                        dev.set_noise_offset(lo - glo)
                        dev.loss_and_grad(x, w, xo, wo, None, beta, lamb, margin, None, sums, grad)
                        _dist.all_reduce_sum(grad)
                        dev.apply_update(grad, optimizer.learning_rate)
                    n_seen += min((batch_idx + 1) * stager.batch_size, stager.n) - batch_idx * stager.batch_size
                    last = batch_idx + 1 == stager.n_batches
                    if sum_batches % 10 == 0 or last:                    # models.py:191-200
                        s = sums.clone()
                        if world > 1: _dist.all_reduce_sum(s)
                        mean = (s / n_seen).cpu().numpy()
                        losses = {'MSE': mean[0]}
                        if beta != 0: losses['KLD'] = mean[1]
                        if lamb != 0: losses['OE'] = mean[2]
                        losses['Train loss'] = mean[3]
                        print('Batch %4d/%d: mean losses' % (sum_batches, total_batches), end='  -->  ', flush=True, file=out)
                        for key, val in losses.items():
                            cond = train_idx + 1 != train_len or not last
                            end = '\r' if cond and key == 'Train loss' else '  '
                            print(key + ' = ' + format(val, '4.3e'), end=end, flush=True, file=out)
            finally:
                stager.close()
        n_valid = 0
        for valid_data in valid_sample:                                      # models.py:201-211
            bkg, ood = valid_data
            stager = _BatchStager([_pick_X(bkg), bkg['weights'], _pick_X(ood), ood['weights']], int(1e6), dev.device, (rank, world))
            try:
                for batch_idx, (x, w, xo, wo) in stager:
                    lo, hi, glo = stager.bounds(batch_idx)
                    dev.set_noise_offset(lo - glo)
                    pj = torch.empty((4, x.shape[0]), dtype=torch.float32, device=dev.device)
                    dev.eval_losses(x, w, xo, wo, None, beta, lamb, margin, pj, vsum)
                n_valid += stager.n
            finally:
                stager.close()
        dev.set_noise_offset(0)
        if world > 1: _dist.all_reduce_sum(vsum)
        losses['Valid loss'] = float(vsum[3].item() / max(n_valid, 1))
        print('Valid loss = ' + format(losses['Valid loss'], '4.3e'), end='  ', flush=True, file=out)
        print('(', '\b' + format(time.time() - start_time, '.1f'), '\b' + 's)', file=out)
        for key in history: history[key] += [float(losses[key]) if key in losses else 0]
        if hist_file is not None and rank == 0: pickle.dump(history, open(hist_file, 'wb'))
        if epoch > 0:
            optimizer, count = model_checkpoint(vae, optimizer, history, model_out, count, patience=patience, _out=out, _save=(rank == 0))
            if count is None: break
        elif n_epochs == 1 and model_out and rank == 0:
            vae.save_weights(model_out)                                      # reference never saves here
    return history


def model_checkpoint(vae, optimizer, history, model_out, count, metric='Train loss', patience=3, factor=2,
                     min_delta=1e-3, min_lr=1e-4, _out=None, _save=True):
    """models.py:222-240.  Returns (optimizer, count); count None stops training."""
    out = _out or sys.stdout
    h = history[metric]
    if h[-1] < np.min(h[:-1]) - min_delta:
        print(metric, 'improved from', format(np.min(h[:-1]), '4.2f'), end=' to ', flush=True, file=out)
        print(format(h[-1], '4.2f'), ' -->  saving model weights to', model_out, file=out)
        if model_out and _save: vae.save_weights(model_out)
        count = 0
    elif h[-1] > np.min(h[-(patience + 1):-1]) - min_delta:
        count += 1
    if count >= patience:
        print('No improvement for', count, 'epochs', end='  -->  ', flush=True, file=out)
        if optimizer.learning_rate < min_lr:
            print('terminating training', file=out)
            count = None
        else:
            print('reducing learning rate from', optimizer.learning_rate, end=' to ', flush=True, file=out)
            optimizer.learning_rate = optimizer.learning_rate / factor
            print(optimizer.learning_rate, file=out)
            count = 0
    return optimizer, count
