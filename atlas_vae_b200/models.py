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
_PINNED = {}                                                  # (address, bytes) -> [array, refcount]: arrays page-locked in place


def _pin(a):
    """cudaHostRegister the array's memory once (a load is stepped through every epoch); _unpin releases it."""
    if a is None or a.nbytes == 0:
        return
    key = (a.ctypes.data, a.nbytes)
    if key in _PINNED:
        _PINNED[key][1] += 1
    elif _ffi.lib().oevae_host_register(a.ctypes.data, a.nbytes) == 0:
        _PINNED[key] = [a, 1]


def _unpin(a):
    if a is None:
        return
    key = (a.ctypes.data, a.nbytes)
    ent = _PINNED.get(key)
    if ent is None:
        return
    ent[1] -= 1
    if ent[1] <= 0:
        _ffi.lib().oevae_host_unregister(a.ctypes.data)
        del _PINNED[key]


def _host_x(a):
    """Constituent rows as the device path takes them: float16 stays float16 (the reference's on-disk dtype, tools/merging.py:31 --
    half the upload; widened on the device), everything else becomes contiguous float32 (Keras casts inputs to float32)."""
    a = np.asarray(a)
    return np.ascontiguousarray(a) if a.dtype == np.float16 else np.ascontiguousarray(a, np.float32)


def _scaler_dev(scaler, D, device):
    """(center, scale) device float64 tensors of an affine scaler (utils.py:406-417), (None, None) without a scaler."""
    if scaler is None:
        return None, None
    from .utils import _scaler_affine
    aff = _scaler_affine(scaler, D)
    if aff is None:
        raise NotImplementedError('train_model(scaler=...) takes affine scalers (MaxAbs / Robust); apply %s with utils.apply_scaler first' % type(scaler).__name__)
    c, sc = aff
    if len(sc) != D or (c is not None and len(c) != D):
        raise ValueError('scaler was fitted on %d columns, the model input has %d' % (len(sc), D))
    return (None if c is None else torch.as_tensor(c).to(device)), torch.as_tensor(sc).to(device)


def _upload(dev, a, center=None, scale=None, stream=None):
    """One host block -> float32 device tensor, widening float16 / applying the affine scaler on the device."""
    t = torch.from_numpy(a).to(dev.device, non_blocking=True)
    if a.dtype == np.float16 or (scale is not None and a.ndim == 2):
        out = torch.empty(t.shape, dtype=torch.float32, device=dev.device)
        if t.numel():
            lib, st = _ffi.lib(), torch.cuda.current_stream().cuda_stream
            if a.dtype == np.float16:
                _ffi.check(lib.oevae_ingest_f16(_ffi.ptr(t), _ffi.ptr(out), t.shape[0], t.shape[1], _ffi.ptr(center), _ffi.ptr(scale), dev.device_index, st))
            else:
                _ffi.check(lib.oevae_affine(_ffi.ptr(t), _ffi.ptr(out), t.shape[0], t.shape[1], _ffi.ptr(center), _ffi.ptr(scale), dev.device_index, st))
        return out
    return t


class _BatchStager:
    """Host (NumPy) load -> device batches for the paths that step from Python (validation chunks; training on the all-reduce
    fallback).  The arrays of a load are page-locked in place once and every batch is copied host->device on a side stream, two
    batches ahead of the compute stream; float16 rows are widened (and an affine scaler applied) on the device."""

    def __init__(self, arrays, batch_size, dev, shard=(0, 1), center=None, scale=None):
        self.dev = dev
        self.device = dev.device
        self.arrays = [None if a is None else (_host_x(a) if np.ndim(a) == 2 else np.ascontiguousarray(a, np.float32)) for a in arrays]
        self.n = len(self.arrays[0])
        self.batch_size = int(batch_size)
        self.n_batches = int(np.ceil(self.n / self.batch_size))
        self.rank, self.world = shard
        self.center, self.scale = center, scale
        self.copy_stream = torch.cuda.Stream(device=self.device)
        for a in self.arrays: _pin(a)
        self.h2d_bytes = 0

    def bounds(self, b):
        lo, hi = b * self.batch_size, min((b + 1) * self.batch_size, self.n)
        if self.world > 1:                      # contiguous shard of the global batch; may be empty in a short last batch
            n = hi - lo
            s0, s1 = lo + (n * self.rank) // self.world, lo + (n * (self.rank + 1)) // self.world
            return s0, s1, lo
        return lo, hi, lo

    def _issue(self, b):
        lo, hi, _ = self.bounds(b)
        out = []
        with torch.cuda.stream(self.copy_stream):
            for a in self.arrays:
                if a is None: out.append(None); continue
                out.append(_upload(self.dev, a[lo:hi], self.center, self.scale))
                self.h2d_bytes += a[lo:hi].nbytes
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
                if t is not None: t.record_stream(torch.cuda.current_stream())
            yield b, tensors

    def close(self):
        torch.cuda.synchronize(self.device)
        for a in self.arrays: _unpin(a)
        self.arrays = []


class _TrainLoad:
    """One (bkg, OoD) load of train_model on the native batch loop (oevae_train_batches_host): page-locked host arrays, optional
    device-resident float32 copy that later epochs step on without another upload (generator OFF: the reference also loads once,
    models.py:139-148)."""

    def __init__(self, dev, bkg, ood, batch_size, shard, center, scale, keep_resident, noise_row0=0):
        self.dev = dev
        self.noise_row0 = int(noise_row0)
        self.X, self.w = _host_x(_pick_X(bkg)), np.ascontiguousarray(bkg['weights'], np.float32)
        self.Xo = self.wo = None
        if ood is not None:
            self.Xo, self.wo = _host_x(_pick_X(ood)), np.ascontiguousarray(ood['weights'], np.float32)
            if self.Xo.dtype != self.X.dtype: self.Xo = self.Xo.astype(self.X.dtype)
            if len(self.Xo) != len(self.X): raise ValueError('OoD sample is not row-aligned with the background sample (utils.py:227)')
        self.n = len(self.X)
        self.batch_size = int(batch_size)
        self.n_batches = int(np.ceil(self.n / self.batch_size))
        self.rank, self.world = shard
        self.center, self.scale = center, scale
        for a in (self.X, self.w, self.Xo, self.wo): _pin(a)
        self.res = None
        D = self.X.shape[1]
        self.stride = -(-self.batch_size // self.world) + 1
        if keep_resident:
            rows = self.stride * self.n_batches
            need = rows * (D + 1) * 4 * (2 if ood is not None else 1)
            free, _ = torch.cuda.mem_get_info(dev.device)
            if need < 0.6 * free:
                mk = lambda *sh: torch.empty(sh, dtype=torch.float32, device=dev.device)
                self.res = [mk(rows, D), mk(rows), mk(rows, D) if ood is not None else None, mk(rows) if ood is not None else None]
        self.filled = False                                  # resident copy complete (after the first pass over the load)
        self.h2d_bytes = 0

    def shard_rows(self, b):
        lo, hi = b * self.batch_size, min((b + 1) * self.batch_size, self.n)
        n = hi - lo
        return (n * self.rank) // self.world, (n * (self.rank + 1)) // self.world

    def run(self, first, count, lr, beta, lamb, margin, sums, peer, trace=None):
        """Steps batches [first, first + count): native host loop (first epoch) or the resident device batches (later epochs)."""
        dev = self.dev
        if self.filled:
            for b in range(first, first + count):
                s0, s1 = self.shard_rows(b)
                o = b * self.stride
                x, w = self.res[0][o:o + s1 - s0], self.res[1][o:o + s1 - s0]
                xo = wo = None
                if self.res[2] is not None: xo, wo = self.res[2][o:o + s1 - s0], self.res[3][o:o + s1 - s0]
                dev.set_noise_offset(self.noise_row0 + s0)
                if peer:
                    dev.train_step_dp(x, w, xo, wo, None, lr, beta, lamb, margin, None, sums)
                else:
                    dev.train_step(x, w, xo, wo, None, lr, beta, lamb, margin, None, sums)
                if trace is not None: trace[b - first].copy_(sums, non_blocking=True)
            dev.set_noise_offset(0)
            return
        L = _ffi.HostLoad()
        L.x_host, L.w_host = self.X.ctypes.data, self.w.ctypes.data
        L.xo_host = None if self.Xo is None else self.Xo.ctypes.data
        L.wo_host = None if self.wo is None else self.wo.ctypes.data
        L.x_dtype = 1 if self.X.dtype == np.float16 else 0
        L.update = 1 if peer else 0
        L.n_rows, L.batch_size, L.first_batch, L.n_batches = self.n, self.batch_size, first, count
        L.rank, L.world, L.noise_row0 = self.rank, self.world, self.noise_row0
        L.center, L.scale = _ffi.ptr(self.center), _ffi.ptr(self.scale)
        if self.res is not None:
            L.x_resident, L.w_resident = _ffi.ptr(self.res[0]), _ffi.ptr(self.res[1])
            L.xo_resident, L.wo_resident = _ffi.ptr(self.res[2]), _ffi.ptr(self.res[3])
            L.resident_stride = self.stride
        L.lr, L.beta, L.lamb, L.margin = lr, beta, lamb, margin
        L.sums_acc = _ffi.ptr(sums)
        L.sums_trace_host = None if trace is None else trace.data_ptr()
        self.h2d_bytes += dev.train_batches_host(L)
        if self.res is not None and first + count >= self.n_batches: self.filled = True

    def close(self):
        torch.cuda.synchronize(self.dev.device)
        for a in (self.X, self.w, self.Xo, self.wo): _unpin(a)
        self.res = None


# ------------------------------------------------------------------------------------------------ train_model
def train_model(vae, train_sample, valid_sample, OE_type='KLD', n_epochs=1, batch_size=5000,
                beta=0, lamb=0, margin=0, lr=1e-3, hist_file=None, model_in=None, model_out=None,
                patience=3, verbose=True, *, scaler=None, metrics_every=10, stats=None, sharded_input=False):
    """models.py:135-219.  Same control flow, history schema, print cadence and LR-patience logic; the batch loop
    (get_losses + tape.gradient + clip + Adam + metric updates, models.py:177-190) runs inside the library: one C-ABI call
    per group of `metrics_every` batches copies the group's batches host->device two ahead of the compute stream and steps them.
    Differences from the reference that do not change results:
      * lamb == 0: the OoD arrays are neither uploaded nor evaluated (the reference multiplies their loss by 0, models.py:130);
      * generator OFF (one load): the load is uploaded once and later epochs step on the device-resident copy;
      * `scaler` (keyword, affine sklearn scaler): the samples are given UNSCALED -- float16 as the reference's files store them
        (tools/merging.py:31) or float32 -- and apply_scaler (utils.py:406-430) runs on the device per batch, bit-identical to the
        host transform and half the upload for float16;
      * `metrics_every` (keyword, default the reference's 10): cadence of the running-mean read-back / print (models.py:191).
    Under torchrun (WORLD_SIZE>1) every global batch is split contiguously across ranks and the gradients are SUMMED before the
    update: inside the update kernel over NVLink peer memory, else by an all-reduce of the flat gradient.
    `sharded_input` (keyword): under torchrun every rank passes ITS OWN part of the sample (equal sizes on all ranks) instead of
    the whole sample: a global batch is then `world x batch_size` jets, `batch_size` from each rank's host (jet-ID's
    `batch_size * n_gpus`, classifier.py:137).  `stats` (keyword): optional dict that receives {'h2d_bytes', 'steps'}."""
    dev = vae.dev
    dev.set_oe_type(OE_type)
    dev.reset_opt()                                           # models.py:138: a new Adam (m = v = 0, t = 0) per train_model call
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
    native = (world == 1 or peer) and dev.noise != 'external'
    own = bool(sharded_input) and world > 1                   # every rank steps through its own rows; sums / counts are global
    shard = (0, 1) if own else (rank, world)
    gmul = world if own else 1
    use_ood = lamb != 0
    center, scale = _scaler_dev(scaler, vae.input_dim, dev.device)
    metrics_every = max(1, int(metrics_every))
    # models.py:191 reads and prints the running means every 10 batches.  With verbose=False nothing is printed, so the read (a stream
    # synchronisation; under torchrun an all-reduce as well) happens once per load and the whole load is ONE native call.  The per-step
    # device->host copy of the running sums (models.py:187-190) is made either way.
    group = metrics_every if verbose else (1 << 30)
    trace = getattr(dev, '_trace', None)                        # page-locked [steps of a group][4] float64, kept on the handle
    total_batches = 0; count = 0; n_steps = 0; h2d = 0
    beta, lamb, margin = float(beta), float(lamb), float(margin)
    loads = {}                                                # generator OFF: the load (and its resident copy) lives across epochs
    vcache = {}
    vready = {}                                               # validation load uploaded while the last training steps of the epoch still run

    def upload_valid(vidx):
        nonlocal h2d
        bkg, ood = valid_sample[vidx]
        arrays = [_pick_X(bkg), bkg['weights']] + ([_pick_X(ood), ood['weights']] if use_ood else [None, None])
        stager = _BatchStager(arrays, int(1e6), dev, shard, center, scale)
        chunks = []
        try:
            for batch_idx, ts in stager:
                lo, hi, glo = stager.bounds(batch_idx)
                chunks.append((lo - glo, ts))
            nv = stager.n
        finally:
            h2d += stager.h2d_bytes
            stager.close()
        chunks = (nv, chunks)
        free, _ = torch.cuda.mem_get_info(dev.device)
        if valid_len == 1 and n_epochs > 1 and free > (8 << 30): vcache[vidx] = chunks
        return chunks
    try:
        for epoch in range(n_epochs):
            sum_batches = 0; n_seen = 0; start_time = time.time()
            print('\nEpoch %d/%d:' % (epoch + 1, n_epochs), file=out)
            sums.zero_(); vsum.zero_()
            losses = {}

            def report(last_of_load, last_load, trace_row=None):
                if trace_row is not None and world == 1:                   # the step's sums already came back (per-step device->host copy)
                    torch.cuda.current_stream().synchronize()
                    mean = trace_row.numpy() / max(n_seen, 1)
                else:
                    s = sums.clone()
                    if world > 1: _dist.all_reduce_sum(s)
                    mean = (s / max(n_seen, 1)).cpu().numpy()
                losses.clear(); losses['MSE'] = mean[0]
                if beta != 0: losses['KLD'] = mean[1]
                if lamb != 0: losses['OE'] = mean[2]
                losses['Train loss'] = mean[3]
                print('Batch %4d/%d: mean losses' % (sum_batches, total_batches), end='  -->  ', flush=True, file=out)
                for key, val in losses.items():
                    cond = (not last_load) or not last_of_load
                    end = '\r' if cond and key == 'Train loss' else '  '
                    print(key + ' = ' + format(val, '4.3e'), end=end, flush=True, file=out)

            for train_idx in range(len(train_sample)):
                last_load = train_idx + 1 == train_len
                if native:
                    ld = loads.get(train_idx)
                    if ld is None:
                        bkg, ood = train_sample[train_idx]
                        ld = _TrainLoad(dev, bkg, ood if use_ood else None, batch_size, shard, center, scale,
                                        keep_resident=(train_len == 1 and n_epochs > 1), noise_row0=rank * int(batch_size) if own else 0)
                        if train_len == 1: loads[train_idx] = ld
                    if epoch == 0: total_batches += ld.n_batches
                    if trace is None or trace.shape[0] < min(group, ld.n_batches):
                        trace = dev._trace = torch.zeros((min(group, max(ld.n_batches, 16)), 4), dtype=torch.float64).pin_memory()
                    try:
                        b = 0
                        while b < ld.n_batches:
                            cnt = min(group - sum_batches % group, ld.n_batches - b)
                            ld.run(b, cnt, optimizer.learning_rate, beta, lamb, margin, sums, peer, trace)
                            b += cnt; sum_batches += cnt; n_steps += cnt
                            n_seen += gmul * (min(b * ld.batch_size, ld.n) - (b - cnt) * ld.batch_size)
                            last = b == ld.n_batches
                            if last and last_load and valid_len == 1 and 0 not in vcache:
                                # the epoch's steps are all enqueued: upload the validation load now, under them, instead of after the
                                # read of the running means below (a stream synchronisation)
                                vready[0] = upload_valid(0)
                            if sum_batches % group == 0 or last:                     # models.py:191-200
                                report(last, last_load, trace[cnt - 1])
                    finally:
                        if train_len != 1: h2d += ld.h2d_bytes; ld.close()
                    continue
                # ---- stepping from Python: all-reduce fallback (gloo / nets too large for the peer-memory update) and EXTERNAL noise
                bkg, ood = train_sample[train_idx]
                arrays = [_pick_X(bkg), bkg['weights']] + ([_pick_X(ood), ood['weights']] if use_ood else [None, None])
                stager = _BatchStager(arrays, batch_size, dev, shard, center, scale)
                if epoch == 0: total_batches += stager.n_batches
                try:
                    for batch_idx, (x, w, xo, wo) in stager:
                        sum_batches += 1; n_steps += 1
                        lo, hi, glo = stager.bounds(batch_idx)
                        dev.set_noise_offset(lo - glo + (rank * int(batch_size) if own else 0))
                        if world == 1:
                            dev.train_step(x, w, xo, wo, None, optimizer.learning_rate, beta, lamb, margin, None, sums)
                        elif peer:
                            dev.train_step_dp(x, w, xo, wo, None, optimizer.learning_rate, beta, lamb, margin, None, sums)
                        else:
                            dev.loss_and_grad(x, w, xo, wo, None, beta, lamb, margin, None, sums, grad)   # empty shard: zero gradient
                            _dist.all_reduce_sum(grad)
                            dev.apply_update(grad, optimizer.learning_rate)
                        n_seen += gmul * (min((batch_idx + 1) * stager.batch_size, stager.n) - batch_idx * stager.batch_size)
                        last = batch_idx + 1 == stager.n_batches
                        if sum_batches % group == 0 or last:                         # models.py:191-200
                            report(last, last_load)
                finally:
                    h2d += stager.h2d_bytes
                    stager.close()
            n_valid = 0
            for vidx, valid_data in enumerate(valid_sample):                         # models.py:201-211
                chunks = vcache.get(vidx) or vready.pop(vidx, None)
                if chunks is None: chunks = upload_valid(vidx)
                nv, parts = chunks
                for off, (x, w, xo, wo) in parts:
                    if x.shape[0] == 0: continue
                    dev.set_noise_offset(off)
                    pj = torch.empty((4, x.shape[0]), dtype=torch.float32, device=dev.device)
                    dev.eval_losses(x, w, xo, wo, None, beta, lamb, margin, pj, vsum)
                n_valid += gmul * nv
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
    finally:
        for ld in loads.values():
            h2d += ld.h2d_bytes
            ld.close()
        if stats is not None:
            stats['h2d_bytes'] = h2d; stats['steps'] = n_steps
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
