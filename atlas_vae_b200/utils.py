"""Hot-path parts of /root/reference/OE-VAE/utils.py on the device (same names and argument meaning).

  jets_sorting    utils.py:564-571      get_4v          utils.py:300-308
  preprocess      utils.py:89-97,115-119 (load_data's sort / keep n_const / zero-pad / n_dims slice)
  fit_scaler      utils.py:389-405 (host sklearn fit, unchanged)   apply_scaler   utils.py:406-430
  loss_function   utils.py:346-375      latent_loss     utils.py:378-386      filtering   utils.py:532-538
  get_idx         utils.py:541-548
  QuantileTransformer / PowerTransformer apply (utils.py:393-396,411) on the device: scaler_kernels.cuh
  OoD_pairing     utils.py:193-227 on the device (m-sort + windowed uniform draw)      OoD_sampling   utils.py:184-190
  reweight_sample / get_weights   utils.py:230-274 (per-load host statistics on <= 10^3 bins, NumPy as in the reference)
HDF5 loading (utils.py:36-130) needs h5py, which this image does not have.
"""
import pickle

import numpy as np
import torch

from . import _ffi
from ._ffi import check, ptr
from .models import HostArray, _host


def _dev(device=None):
    if not torch.cuda.is_available():
        raise _ffi.OevaeError('no CUDA device: the OE-VAE hot path has no CPU fallback')
    return torch.device('cuda', torch.cuda.current_device() if device is None else device)


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _to_dev(a, dev, dtype=torch.float32):
    if isinstance(a, torch.Tensor):
        return a.to(dev, dtype).contiguous()
    return torch.as_tensor(np.ascontiguousarray(a)).to(dev, dtype).contiguous()


def _scaler_affine(scaler, D):
    """(center, scale) float64 device-ready arrays of an affine sklearn scaler, or None if not affine."""
    name = type(scaler).__name__
    if name == 'MaxAbsScaler':
        return None, np.asarray(scaler.scale_, np.float64)
    if name == 'RobustScaler':
        c = np.asarray(scaler.center_, np.float64) if scaler.with_centering else None
        s = np.asarray(scaler.scale_, np.float64) if scaler.with_scaling else np.ones(D)
        return c, s
    return None


_NDTRI_CLIP = (-5.199337582605575, 5.19933758270342)        # stats.norm.ppf(1e-7 - np.spacing(1)), stats.norm.ppf(1 - (1e-7 - np.spacing(1)))


def transform_device(x_t, scaler, out=None):
    """apply_scaling (utils.py:406-417) of a fitted sklearn scaler on a device float32 [N, D] tensor (in place unless `out`):
    MaxAbs / Robust (affine, bit-exact), QuantileTransformer(normal) (bit-exact), PowerTransformer (yeo-johnson; float32 ulps)."""
    dev = x_t.device
    N, D = x_t.shape
    out = x_t if out is None else out
    lib = _ffi.lib()
    name = type(scaler).__name__
    aff = _scaler_affine(scaler, D)
    if aff is not None:
        c, sc = aff
        if len(sc) != D or (c is not None and len(c) != D):
            raise ValueError('scaler was fitted on %d columns, the array has %d' % (len(sc), D))
        c_t = None if c is None else torch.as_tensor(c).to(dev)
        s_t = torch.as_tensor(sc).to(dev)
        check(lib.oevae_affine(ptr(x_t), ptr(out), N, D, ptr(c_t), ptr(s_t), dev.index, _stream()))
    elif name == 'QuantileTransformer':
        if scaler.output_distribution != 'normal':
            raise NotImplementedError("QuantileTransformer: only output_distribution='normal' (the reference's, utils.py:394) is on the device")
        q = np.ascontiguousarray(np.asarray(scaler.quantiles_, np.float64).T)              # [D, n_q]: a column's table is contiguous
        if q.shape[0] != D: raise ValueError('scaler was fitted on %d columns, the array has %d' % (q.shape[0], D))
        q_t = torch.as_tensor(q).to(dev); r_t = torch.as_tensor(np.asarray(scaler.references_, np.float64)).to(dev)
        check(lib.oevae_quantile_normal(ptr(x_t), ptr(out), N, D, ptr(q_t), ptr(r_t), q.shape[1], _NDTRI_CLIP[0], _NDTRI_CLIP[1], dev.index, _stream()))
    elif name == 'PowerTransformer':
        if scaler.method != 'yeo-johnson':
            raise NotImplementedError("PowerTransformer: only method='yeo-johnson' (the default the reference uses, utils.py:396)")
        lam = np.asarray(scaler.lambdas_)
        if lam.dtype != np.float32:
            raise NotImplementedError('PowerTransformer fitted on float64 data: the reference fits on float32 constituents (utils.py:97,390-400)')
        if len(lam) != D: raise ValueError('scaler was fitted on %d columns, the array has %d' % (len(lam), D))
        l_t = torch.as_tensor(lam).to(dev)
        m_t = s_t = None
        if scaler.standardize:
            st = scaler._scaler
            m_t = torch.as_tensor(np.asarray(st.mean_, np.float32)).to(dev) if st.with_mean else None
            s_t = torch.as_tensor(np.asarray(st.scale_, np.float32)).to(dev) if st.with_std else None
        check(lib.oevae_power_yj(ptr(x_t), ptr(out), N, D, ptr(l_t), ptr(m_t), ptr(s_t), dev.index, _stream()))
    else:
        raise NotImplementedError('%s has no device transform' % name)
    return out


def preprocess_device(raw, n_const, n_dims=4, scaler=None, sort=True, device=None, want_4v=True):
    """K0 on device tensors: raw [N, 4*n_raw] (fp32 or fp16; E,px,py,pz) -> (X [N,n_dims*n_const], pt, m).  Affine scalers are
    folded into the K0 kernel; Quantile / Power transformers run as a second kernel on its output."""
    dev = _dev(device)
    if isinstance(raw, torch.Tensor):
        raw_t = raw.to(dev).contiguous()
    else:
        raw_t = torch.as_tensor(np.ascontiguousarray(raw)).to(dev)
    if raw_t.dtype == torch.float16:
        code = 1
    else:
        raw_t = raw_t.to(torch.float32); code = 0
    N, n_raw = raw_t.shape[0], raw_t.shape[1] // 4
    D = n_dims * n_const
    c_t = s_t = None
    aff = None if scaler is None else _scaler_affine(scaler, D)
    if aff is not None:
        c, s = aff
        if len(s) != D or (c is not None and len(c) != D):
            raise ValueError('scaler was fitted on %d columns, the preprocessed input has %d (fit it with reshape=False on [N, n_dims * n_const])' % (len(s), D))
        c_t = None if c is None else torch.as_tensor(c).to(dev)
        s_t = torch.as_tensor(s).to(dev)
    x = torch.empty((N, D), dtype=torch.float32, device=dev)
    pt = torch.empty(N, dtype=torch.float32, device=dev) if want_4v else None
    m = torch.empty(N, dtype=torch.float32, device=dev) if want_4v else None
    check(_ffi.lib().oevae_preprocess(ptr(raw_t), code, N, n_raw, int(n_const), int(n_dims), ptr(c_t), ptr(s_t),
                                      1 if sort else 0, ptr(x), ptr(pt), ptr(m), dev.index, _stream()))
    if scaler is not None and aff is None:
        transform_device(x, scaler)
    return x, pt, m


def jets_sorting(jets, idx=(0, None), return_dict=None):
    """utils.py:564-571: constituents re-ordered by decreasing pT; same shape as the input, float32."""
    jets = np.asarray(jets)[idx[0]:idx[1]]
    x, _, _ = preprocess_device(jets, jets.shape[1] // 4, 4, None, True, want_4v=False)
    out = _host(x)
    if idx == (0, None): return out
    return_dict[idx] = out


def get_4v(sample, idx=(0, None), return_dict=None):
    """utils.py:300-308: jet pt and mass from the 4-vector sum of the given constituents (stored order)."""
    sample = np.asarray(sample)[idx[0]:idx[1]]
    _, pt, m = preprocess_device(sample, sample.shape[1] // 4, 4, None, False)
    out = {'pt_calo': _host(pt), 'm_calo': _host(m)}
    if idx == (0, None): return out
    return_dict[idx] = out


def preprocess(raw, n_const, n_dims=4, scaler=None):
    """load_data's constituent path (utils.py:89-97,115-119) + get_4v + affine scaler in one kernel.
    Returns {'constituents', 'pt', 'm'} as host arrays."""
    x, pt, m = preprocess_device(raw, n_const, n_dims, scaler, True)
    return {'constituents': _host(x), 'pt': _host(pt), 'm': _host(m)}


def fit_scaler(sample, n_dims, scaler_out, scaler_type='RobustScaler', reshape=False):
    """utils.py:389-405.  The fit is a one-off host statistic (sklearn, as in the reference)."""
    from sklearn import preprocessing
    if reshape: sample = np.reshape(sample, (-1, n_dims))
    if scaler_type == 'QuantileTransformer':
        scaler = preprocessing.QuantileTransformer(output_distribution='normal', n_quantiles=10000, random_state=0)
    elif scaler_type == 'PowerTransformer':
        scaler = preprocessing.PowerTransformer()
    elif scaler_type == 'RobustScaler':
        scaler = preprocessing.RobustScaler()
    elif scaler_type == 'MaxAbsScaler':
        scaler = preprocessing.MaxAbsScaler()
    else:
        raise ValueError(scaler_type)
    scaler.fit(sample)
    if scaler_out:
        pickle.dump(scaler, open(scaler_out, 'wb'))
    return scaler


def apply_scaler(sample, n_dims, scaler, tag='sample', reshape=False, verbose=False):
    """utils.py:419-430 on the device: MaxAbs / Robust (affine), QuantileTransformer(normal) and PowerTransformer(yeo-johnson).
    Returns a float32 host array of the input's shape (sklearn's transforms of a float32 array are float32 too)."""
    if scaler is None: return sample
    shape = np.shape(sample)
    x2 = np.reshape(sample, (-1, n_dims)) if reshape else np.asarray(sample)
    x_t = _to_dev(x2, _dev())
    if x_t.data_ptr() == np.asarray(x2).ctypes.data: x_t = x_t.clone()
    return _host(transform_device(x_t, scaler)).reshape(shape)


_METRIC_SLOT = {'MSE': 0, 'MAE': 1, 'MARE': 2, 'KLD': 3, 'JSD': 4, 'X-S': 5}


def loss_function(P, Q, n_dims, metric, X_losses=None, delta=1e-32, multiloss=True):
    """utils.py:346-375: per-jet metric between X_true (P) and X_pred (Q), computed in float32 on the device."""
    dev = _dev()
    if metric in ('Inputs', 'Inputs_scaled'):
        loss = _host(_to_dev(P, dev).mean(dim=1))
    else:
        if metric not in _METRIC_SLOT: raise ValueError(metric)
        P_t, Q_t = _to_dev(P, dev), _to_dev(Q, dev)
        out = torch.empty(P_t.shape[0], dtype=torch.float32, device=dev)
        args = [None] * 6
        args[_METRIC_SLOT[metric]] = ptr(out)
        check(_ffi.lib().oevae_pair_metrics(ptr(P_t), ptr(Q_t), P_t.shape[0], P_t.shape[1], *args, dev.index, _stream()))
        loss = _host(out)
    if multiloss: X_losses[metric] = loss
    else: return loss


def latent_loss(X_true, model):
    """utils.py:378-386: per-jet latent KLD, non-finite -> 0."""
    out = model.dev.score(X_true, 1, None if model.dev.noise != 'external' else
                          torch.zeros((len(X_true), model.dev.latent_dim), device=model.dev.device),
                          metrics=('Latent',), want_finite=False)
    return _host(out['Latent'])


def filtering(y_true, X_true, X_pred, sample):
    """utils.py:532-538: drop the rows whose prediction has a non-finite value."""
    keep = np.all(np.isfinite(X_pred), axis=1)
    if keep.all(): return y_true, X_true, X_pred, sample
    for key in sample: sample[key] = sample[key][keep]
    return y_true[keep], X_true[keep], X_pred[keep], sample


def get_idx(max_val, n_bins=10, bin_size=None, min_val=0, integer=True, tuples=True):
    """utils.py:541-548 (host helper: bin edges / index ranges)."""
    if bin_size is None:
        n_bins = max(1, min(max_val - min_val, n_bins))
        bin_size = (max_val - min_val) // n_bins
    idx_list = np.append(np.arange(min_val, max_val, bin_size), max_val)
    if integer: idx_list = np.int_(idx_list)
    if tuples: return list(zip(idx_list[:-1], idx_list[1:]))
    return idx_list


# ------------------------------------------------------------------------------------------------ sample assembly (before training)
def OoD_sampling(bkg_sample, OoD_sample, adjust_weights=False, seed=None):
    """utils.py:184-190: resample the OoD sample to the background's size (host NumPy, as in the reference)."""
    np.random.seed(seed)
    source_size = len(list(OoD_sample.values())[0])
    target_size = len(list(bkg_sample.values())[0])
    indices = np.random.choice(source_size, target_size, replace=source_size < target_size)
    if adjust_weights: OoD_sample['weights'] = OoD_sample['weights'] * np.float32(source_size / target_size)
    return {key: np.take(OoD_sample[key], indices, axis=0) for key in OoD_sample}


def OoD_pairing(bkg_sample, OoD_sample, multithread=True, verbose=True, seed=0, pick_first=False, return_indices=False):
    """utils.py:193-227 on the device: the OoD sample is sorted by mass (radix argsort) and every background jet gets one OoD jet drawn
    uniformly from the (m, pT) window around it -- 10 x 10 GeV, doubled alternately in pT and in m until it is not empty (utils.py:
    198-210).  `multithread` is accepted for compatibility (the reference forks 16 processes; here one warp per background jet).
    The draw is Philox4x32-10 keyed by (seed, jet): np.random.choice under fork() cannot be reproduced, the candidate sets are the
    reference's (tests/test_gpu_pairing.py against its own function body).  Returns the paired OoD sample, row-aligned with bkg_sample."""
    import ctypes as C
    import time
    dev = _dev()
    t0 = time.time()
    if verbose: print('Pairing OoD with QCD', end=' ', flush=True)
    m_o, pt_o = _to_dev(OoD_sample['m'], dev), _to_dev(OoD_sample['pt'], dev)
    n_o = m_o.numel()
    order = torch.empty(n_o, dtype=torch.int64, device=dev)
    check(_ffi.lib().oevae_argsort_f32(ptr(m_o), n_o, ptr(order), dev.index, _stream()))
    m_s, pt_s = m_o[order].contiguous(), pt_o[order].contiguous()
    m_b, pt_b = _to_dev(bkg_sample['m'], dev), _to_dev(bkg_sample['pt'], dev)
    n_b = m_b.numel()
    idx = torch.empty(n_b, dtype=torch.int64, device=dev)
    check(_ffi.lib().oevae_ood_pairing(ptr(m_b), ptr(pt_b), n_b, ptr(m_s), ptr(pt_s), n_o, int(seed) & (2 ** 64 - 1), 1 if pick_first else 0,
                                       ptr(idx), None, dev.index, _stream()))
    take = order[idx.clamp_min(0)].cpu().numpy()                            # indices into the caller's (unsorted) OoD arrays
    if bool((idx < 0).any()): raise ValueError('OoD_pairing: empty OoD sample')
    if verbose: print('(', '\b' + format(time.time() - t0, '2.1f'), '\b' + ' s)')
    if return_indices: return take
    return {key: np.take(OoD_sample[key], take, axis=0) for key in OoD_sample}


def get_weights(bkg_sample, sig_sample, bin_sizes, weight_type, max_val=1e4, density=True):
    """utils.py:253-274: (m, pT)-histogram ratio weights.  A per-load host statistic on <= 10^3 bins: NumPy's own histogram2d /
    digitize, as in the reference."""
    m_size, pt_size = bin_sizes['m'], bin_sizes['pt']
    m_bkg, pt_bkg, weights_bkg = [bkg_sample[key] for key in ['m', 'pt', 'weights']]
    m_sig, pt_sig, weights_sig = [sig_sample[key] for key in ['m', 'pt', 'weights']]
    m_min, pt_min = np.min(m_sig), np.min(pt_sig)
    m_max, pt_max = np.max(m_sig), np.max(pt_sig)
    if 'm' in weight_type: pt_size = pt_max + 1
    if 'pt' in weight_type: m_size = m_max + 1
    m_bins = get_idx(m_max, bin_size=m_size, min_val=m_min, integer=False, tuples=False)
    pt_bins = get_idx(pt_max, bin_size=pt_size, min_val=pt_min, integer=False, tuples=False)
    m_idx = np.clip(np.digitize(m_sig, m_bins, right=False), 1, len(m_bins) - 1) - 1
    pt_idx = np.clip(np.digitize(pt_sig, pt_bins, right=False), 1, len(pt_bins) - 1) - 1
    hist_sig = np.histogram2d(m_sig, pt_sig, bins=[m_bins, pt_bins], density=density)[0]
    if density: hist_sig *= len(m_sig)
    hist_sig = np.maximum(hist_sig, np.min(hist_sig[hist_sig != 0]) if density else 1)
    if 'flat' in weight_type:
        weights = (1 / hist_sig)[m_idx, pt_idx]
        return weights * np.sum(weights_sig) / np.sum(weights)
    hist_bkg = np.histogram2d(m_bkg, pt_bkg, bins=[m_bins, pt_bins], weights=weights_bkg, density=density)[0]
    if density: hist_bkg *= len(m_bkg)
    weights = (hist_bkg / hist_sig)[m_idx, pt_idx]
    return np.minimum(max_val, weights * np.sum(weights_bkg) / np.sum(weights))


def reweight_sample(bkg_sample, sig_sample, bin_sizes, weight_type='X-S'):
    """utils.py:230-250: None / X-S / flat_m / flat_pt / flat_2d / OoD_m / OoD_pt / OoD_2d (same branches, same order)."""
    if weight_type is None or weight_type.lower() == 'none':
        sig_sample['weights'] = np.ones_like(sig_sample['weights'])
        bkg_sample['weights'] = np.ones_like(bkg_sample['weights'])
        weight_type = 'none'
    if 'flat' in weight_type:
        sig_sample['weights'] = get_weights(bkg_sample, sig_sample, bin_sizes, weight_type)
        bkg_sample['weights'] = get_weights(bkg_sample, bkg_sample, bin_sizes, weight_type)
        sig_sample['weights'] = get_weights(bkg_sample, sig_sample, bin_sizes, weight_type='2d')
    if 'OoD' in weight_type:
        sig_sample['weights'] = get_weights(bkg_sample, sig_sample, bin_sizes, weight_type)
    if weight_type == 'X-S':
        sig_sample['weights'] *= np.sum(bkg_sample['weights']) / np.sum(sig_sample['weights'])
    return bkg_sample, sig_sample
