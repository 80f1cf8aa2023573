"""Thin object wrapper over an oevae_t handle: device tensors in, device tensors out, no host sync.

Everything numerical happens in liboevae.so; torch only allocates device memory and supplies the stream.
"""
import ctypes as C
import numpy as np
import torch

from . import _ffi
from ._ffi import check, ptr, OevaeError


def _stream():
    return torch.cuda.current_stream().cuda_stream


class DeviceVAE:
    """One handle per (process, GPU).  FC_layers = hidden widths + [latent_dim] (models.py:22,27-28)."""

    def __init__(self, FC_layers, input_dim, OE_type='KLD', precision='fp32', noise='fixed', seed=0,
                 device=None, max_batch=0):
        if not torch.cuda.is_available():
            raise OevaeError('no CUDA device: the OE-VAE hot path has no CPU fallback')
        self.lib = _ffi.lib()
        self.FC_layers = [int(v) for v in FC_layers]
        self.input_dim = int(input_dim)
        self.latent_dim = self.FC_layers[-1]
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.device = torch.device('cuda', self.device_index)
        cfg = _ffi.Config()
        cfg.input_dim = self.input_dim
        cfg.n_hidden = len(self.FC_layers) - 1
        for i, h in enumerate(self.FC_layers[:-1]):
            cfg.hidden[i] = h
        cfg.latent_dim = self.latent_dim
        cfg.oe_type = _ffi.OE_TYPES[OE_type]
        cfg.precision = _ffi.PRECISIONS[precision]
        cfg.noise_mode = _ffi.NOISE_MODES[noise]
        cfg.device = self.device_index
        cfg.seed = int(seed) & (2 ** 64 - 1)
        cfg.max_batch = int(max_batch)
        self.noise = noise
        self.OE_type = OE_type
        h = C.c_void_p()
        rc = self.lib.oevae_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise OevaeError('oevae_create failed (%d): %s' % (rc, self.lib.oevae_last_error(None).decode()))
        self.h = h
        self.n_params = int(self.lib.oevae_n_params(self.h))
        self.dp = False                                       # peer-memory data-parallel path connected (dp_connect)

    def close(self):
        if getattr(self, 'h', None):
            self.lib.oevae_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- layout / weights
    def layout(self):
        n = 2 * (len(self.FC_layers) - 1) + 3
        off = (C.c_int64 * (2 * n))(); di = (C.c_int32 * n)(); do = (C.c_int32 * n)()
        got = self.lib.oevae_param_layout(self.h, off, di, do, n)
        assert got == n
        return [(int(off[2 * i]), int(off[2 * i + 1]), int(di[i]), int(do[i])) for i in range(n)]

    def set_oe_type(self, OE_type):
        if OE_type != self.OE_type:
            check(self.lib.oevae_set_oe_type(self.h, _ffi.OE_TYPES[OE_type]), self.h)
            self.OE_type = OE_type

    def set_params(self, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        check(self.lib.oevae_set_params_host(self.h, ptr(flat), flat.size), self.h)

    def get_params(self):
        flat = np.empty(self.n_params, np.float32)
        check(self.lib.oevae_get_params_host(self.h, ptr(flat), flat.size), self.h)
        return flat

    def set_opt_state(self, m, v, step):
        m = np.ascontiguousarray(m, np.float32); v = np.ascontiguousarray(v, np.float32)
        check(self.lib.oevae_set_opt_state_host(self.h, ptr(m), ptr(v), m.size, int(step)), self.h)

    def get_opt_state(self):
        m = np.empty(self.n_params, np.float32); v = np.empty(self.n_params, np.float32); st = C.c_int64()
        check(self.lib.oevae_get_opt_state_host(self.h, ptr(m), ptr(v), m.size, C.byref(st)), self.h)
        return m, v, int(st.value)

    def reset_opt(self):
        """Fresh Adam (m = v = 0, step = 0): what `tf.keras.optimizers.Adam(lr)` at models.py:138 gives every train_model call."""
        check(self.lib.oevae_reset_opt_state(self.h), self.h)

    def dp_last_grad(self):
        """SUM over ranks of the last train_step_dp's gradient, before clip_values (host array)."""
        g = np.empty(self.n_params, np.float32)
        check(self.lib.oevae_dp_last_grad_host(self.h, ptr(g), g.size), self.h)
        return g

    def set_noise_offset(self, row_offset):
        check(self.lib.oevae_set_noise_offset(self.h, int(row_offset)), self.h)

    # ---- helpers
    def _f32(self, t):
        if t is None:
            return None
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(np.ascontiguousarray(t, np.float32))
        if t.device != self.device or t.dtype != torch.float32 or not t.is_contiguous():
            t = t.to(self.device, torch.float32).contiguous()
        return t

    def _eps(self, eps, B):
        if eps is None:
            return None, 0
        eps = self._f32(eps)
        n_eps = 1 if eps.dim() == 2 else eps.shape[0]
        assert eps.shape[-1] == self.latent_dim and eps.shape[-2] == B, 'eps must be [B,L] or [n,B,L]'
        return eps, n_eps

    # ---- training / evaluation
    def train_step(self, x, w, x_ood=None, w_ood=None, eps=None, lr=1e-3, beta=0., lamb=0., margin=0.,
                   per_jet=None, sums=None):
        x, w, x_ood, w_ood = self._f32(x), self._f32(w), self._f32(x_ood), self._f32(w_ood)
        B = x.shape[0]
        eps, n_eps = self._eps(eps, B)
        check(self.lib.oevae_train_step(self.h, ptr(x), ptr(w), ptr(x_ood), ptr(w_ood), ptr(eps), n_eps, B,
                                        lr, beta, lamb, margin, ptr(per_jet), ptr(sums), _stream()), self.h)

    def dp_connect(self):
        """Data parallel over NVLink peer memory: exchange the IPC handles of the per-rank gradient buffers (torch.distributed
        is only the transport of 64 bytes per rank).  Returns False when the fast path does not apply (then use
        loss_and_grad + all_reduce + apply_update)."""
        import torch.distributed as td
        if not (td.is_available() and td.is_initialized() and td.get_world_size() > 1):
            return False
        rank, world = td.get_rank(), td.get_world_size()
        ok = world <= 8
        raw = C.create_string_buffer(64)
        if ok:
            check(self.lib.oevae_dp_export(self.h, raw), self.h)
        handles = [None] * world
        td.all_gather_object(handles, raw.raw if ok else None)
        if any(h is None for h in handles):
            return False
        blob = C.create_string_buffer(b''.join(handles), 64 * world)
        rc = self.lib.oevae_dp_connect(self.h, rank, world, blob)
        flags = [None] * world
        td.all_gather_object(flags, int(rc))
        self.dp = all(f == 0 for f in flags)                 # every rank takes the same path
        td.barrier()
        return self.dp

    def train_step_dp(self, x, w, x_ood=None, w_ood=None, eps=None, lr=1e-3, beta=0., lamb=0., margin=0.,
                      per_jet=None, sums=None):
        x, w, x_ood, w_ood = self._f32(x), self._f32(w), self._f32(x_ood), self._f32(w_ood)
        B = x.shape[0]
        eps, n_eps = self._eps(eps, B)
        check(self.lib.oevae_train_step_dp(self.h, ptr(x), ptr(w), ptr(x_ood), ptr(w_ood), ptr(eps), n_eps, B,
                                           lr, beta, lamb, margin, ptr(per_jet), ptr(sums), _stream()), self.h)

    def train_batches_host(self, load):
        """oevae_train_batches_host: `load` is an _ffi.HostLoad; returns the host->device bytes the call copied."""
        check(self.lib.oevae_train_batches_host(self.h, C.byref(load), _stream()), self.h)
        return int(load.h2d_bytes)

    def loss_and_grad(self, x, w, x_ood=None, w_ood=None, eps=None, beta=0., lamb=0., margin=0.,
                      per_jet=None, sums=None, grad=None):
        x, w, x_ood, w_ood = self._f32(x), self._f32(w), self._f32(x_ood), self._f32(w_ood)
        B = x.shape[0]
        eps, n_eps = self._eps(eps, B)
        if grad is None:
            grad = torch.empty(self.n_params, dtype=torch.float32, device=self.device)
        check(self.lib.oevae_loss_and_grad(self.h, ptr(x), ptr(w), ptr(x_ood), ptr(w_ood), ptr(eps), n_eps, B,
                                           beta, lamb, margin, ptr(per_jet), ptr(sums), ptr(grad), _stream()), self.h)
        return grad

    def apply_update(self, grad, lr):
        check(self.lib.oevae_apply_update(self.h, ptr(grad), lr, _stream()), self.h)

    def eval_losses(self, x, w, x_ood=None, w_ood=None, eps=None, beta=0., lamb=0., margin=0.,
                    per_jet=None, sums=None):
        x, w, x_ood, w_ood = self._f32(x), self._f32(w), self._f32(x_ood), self._f32(w_ood)
        B = x.shape[0]
        eps, n_eps = self._eps(eps, B)
        if per_jet is None:
            per_jet = torch.empty((4, B), dtype=torch.float32, device=self.device)
        check(self.lib.oevae_eval_losses(self.h, ptr(x), ptr(w), ptr(x_ood), ptr(w_ood), ptr(eps), n_eps, B,
                                         beta, lamb, margin, ptr(per_jet), ptr(sums), _stream()), self.h)
        return per_jet

    def gemm(self, kind, A, B, bias=None, relu=False):
        """Kernel-level access to the layered contraction (tests): kind 0: A[R,T] @ B[T,C]; 1: A[R,T] @ B[C,T]^T;
        2: A[T,R]^T @ B[T,C]."""
        A, B, bias = self._f32(A), self._f32(B), self._f32(bias)
        if kind == 0: R, T, Cn = A.shape[0], A.shape[1], B.shape[1]; st = (A.shape[1], 1, B.shape[1], 1)
        elif kind == 1: R, T, Cn = A.shape[0], A.shape[1], B.shape[0]; st = (A.shape[1], 1, 1, B.shape[1])
        else: R, T, Cn = A.shape[1], A.shape[0], B.shape[1]; st = (1, A.shape[1], B.shape[1], 1)
        out = torch.empty((R, Cn), dtype=torch.float32, device=self.device)
        check(self.lib.oevae_gemm(self.h, kind, ptr(A), st[0], st[1], ptr(B), st[2], st[3], ptr(out), Cn, R, Cn, T,
                                  ptr(bias), 1 if relu else 0, _stream()), self.h)
        return out

    def encode(self, x, eps=None):
        x = self._f32(x); N = x.shape[0]
        eps = self._f32(eps)
        mu = torch.empty((N, self.latent_dim), dtype=torch.float32, device=self.device)
        lv = torch.empty_like(mu); z = torch.empty_like(mu)
        want_z = not (self.noise == 'external' and eps is None)
        check(self.lib.oevae_encode(self.h, ptr(x), N, ptr(eps), ptr(mu), ptr(lv), ptr(z) if want_z else None, _stream()), self.h)
        return mu, lv, (z if want_z else None)

    def score(self, x, n_iter=1, eps=None, metrics=('MSE', 'MAE', 'KLD', 'JSD', 'Latent'), want_xhat=False, want_finite=True):
        """Single pass over x: per-jet metrics of utils.loss_function + latent KLD + finite flag."""
        x = self._f32(x); N = x.shape[0]
        eps = self._f32(eps)
        out = {m: torch.empty(N, dtype=torch.float32, device=self.device) for m in metrics}
        fin = torch.empty(N, dtype=torch.uint8, device=self.device) if want_finite else None
        xhat = torch.empty((N, self.input_dim), dtype=torch.float32, device=self.device) if want_xhat else None
        g = lambda k: ptr(out[k]) if k in out else None
        check(self.lib.oevae_score(self.h, ptr(x), N, int(n_iter), ptr(eps), g('MSE'), g('MAE'), g('KLD'), g('JSD'),
                                   g('Latent'), ptr(fin), ptr(xhat), _stream()), self.h)
        if want_finite:
            out['finite'] = fin
        if want_xhat:
            out['xhat'] = xhat
        return out
