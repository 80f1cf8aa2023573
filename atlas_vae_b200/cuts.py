"""Arithmetic of the reference's scoring / cut / histogram chain on the device (no figure drawing).

  score_sample       vae.py:163-167 + plots.py:22-35   (predict x n_iter, loss_function metrics, latent_loss,
                                                        filtering, loss_mapping) in ONE pass over the jets
  loss_mapping       plots.py:15-20
  get_rates          plots.py:228-232   (host sklearn roc_curve: threshold search is SURVEY 8(f) f2)
  best_threshold     plots.py:235-247
  make_cut           plots.py:250-259
  mass_histograms    plots.py:697-715   (bin edges from get_idx, per-class weighted histogram)
  generate_cuts      plots.py:88-104    (the six bkg_eff working points + best gain; returns histograms)
"""
import numpy as np
import torch

from . import _ffi, dist as _dist
from ._ffi import check, ptr
from .models import _host
from .utils import _dev, _stream, _to_dev, get_idx


def loss_mapping_branch(lo, hi, has_nan=False):
    """Which branch of plots.py:15-20 applies, from the global min/max (np.all over comparisons is False
    as soon as one element is NaN, which sends the reference to the last branch)."""
    if has_nan: return 4
    if lo >= 0 and hi <= 1: return 0
    if lo >= -1 and hi <= 0: return 1
    if lo >= 0: return 2
    if hi <= 0: return 3
    return 4


def loss_mapping_device(x_t, keep=None):
    """In-place loss_mapping of a device score vector; min/max are all-reduced across ranks."""
    dev = x_t.device
    mm = torch.tensor([np.inf, -np.inf, 0.0], dtype=torch.float64, device=dev)
    check(_ffi.lib().oevae_minmax(ptr(x_t), ptr(keep), x_t.numel(), ptr(mm), dev.index, _stream()))
    lo, hi, nn = mm[0:1].clone(), mm[1:2].clone(), mm[2:3].clone()
    _dist.all_reduce_minmax(lo, hi); _dist.all_reduce_sum(nn)
    branch = loss_mapping_branch(float(lo), float(hi), float(nn) > 0)
    check(_ffi.lib().oevae_loss_map(ptr(x_t), x_t.numel(), branch, dev.index, _stream()))
    return x_t


def loss_mapping(x):
    """plots.py:15-20 on a host array (device arithmetic)."""
    x_t = _to_dev(np.asarray(x, np.float32), _dev())
    if x_t.data_ptr() == 0 or x_t.numel() == 0: return np.asarray(x, np.float32)
    return _host(loss_mapping_device(x_t.clone()))


def score_sample(model, X_true, n_iter=1, metrics=('Latent', 'MAE', 'KLD', 'JSD'), normal_losses='ON', eps=None,
                 return_device=False):
    """One device pass replacing predict (vae.py:163-166), filtering (utils.py:532-538), the per-metric
    loss_function processes (plots.py:22-29), latent_loss (plots.py:31-32) and loss_mapping (plots.py:34-35).
    Returns ({metric: scores[N]}, finite[N] bool)."""
    out = model.dev.score(X_true, n_iter, eps, metrics=tuple(metrics), want_finite=True)
    fin = out.pop('finite')
    if normal_losses == 'ON':
        for k in out:
            loss_mapping_device(out[k], fin)
    if return_device:
        return out, fin
    return {k: _host(v) for k, v in out.items()}, _host(fin).astype(bool)


def get_rates(y_true, X_loss, weights, metric=None, return_dict=None):
    """plots.py:228-232 (weighted ROC by sklearn, as in the reference)."""
    from sklearn import metrics
    fpr, tpr, thresholds = metrics.roc_curve(y_true, X_loss, pos_label=0, sample_weight=weights)
    sel = fpr != 0
    res = 100 * fpr[sel], 100 * tpr[sel], thresholds[sel]
    if return_dict is None: return res
    return_dict[metric] = res


def best_threshold(y_true, positive_rates, weights, cut_type, min_tpr=1):
    """plots.py:235-247"""
    fpr, tpr, thresholds = positive_rates
    sel = tpr > min_tpr
    fpr, tpr, thresholds = fpr[sel], tpr[sel], thresholds[sel]
    if cut_type == 'gain':
        cut_values = tpr / fpr; factor = 1
    elif cut_type == 'sigma':
        cut_values = tpr / np.sqrt(fpr)
        factor = np.sum(weights[y_true == 0]) / np.sqrt(np.sum(weights[y_true == 1])) / 10
    else:
        raise ValueError(cut_type)
    i = np.argmax(cut_values)
    return thresholds[i], cut_values[i] * factor


def cut_value(y_true, positive_rates, weights, cut_type='gain', bkg_eff=None):
    """Threshold half of make_cut (plots.py:250-258)."""
    if bkg_eff is None:
        return best_threshold(y_true, positive_rates, weights, cut_type)[0]
    fpr, tpr, thresholds = positive_rates
    return thresholds[np.argmin(abs(fpr - bkg_eff))]


def make_cut(y_true, X_loss, sample, positive_rates, metric, cut_type, bkg_eff=None):
    """plots.py:250-259: {key: sample[key][X_loss > loss_cut]}"""
    loss_cut = cut_value(y_true, positive_rates, sample['weights'], cut_type, bkg_eff)
    mask = np.asarray(X_loss) > loss_cut
    return {key: sample[key][mask] for key in sample}


def bin_edges(var, bin_size):
    """plots.py:697-700: get_idx(max, bin_size, min_val=max(0,min), integer=False, tuples=False) -> float64 edges."""
    return get_idx(np.max(var), bin_size=bin_size, min_val=max(0, np.min(var)), integer=False, tuples=False)


def mass_histograms_device(var_t, w_t, jzw_t, edges, score_t=None, cut=0.0, keep_t=None, reduce=True):
    """K5: per-class (row 0: JZW==-1, row 1: JZW>=0) histogram of `var` for the jets with score > cut.
    Returns (counts int64 [2,nb], wsums float64 [2,nb]) on the host; summed over ranks when `reduce`."""
    dev = var_t.device
    edges_t = torch.as_tensor(np.ascontiguousarray(edges, np.float64)).to(dev)
    nb = edges_t.numel() - 1
    if nb < 1:                                  # a selection with min == max gives a single edge (utils.py:541-548): no bins
        return np.zeros((2, 0), np.int64), np.zeros((2, 0), np.float64)
    counts = torch.zeros((2, nb), dtype=torch.int64, device=dev)
    wsums = torch.zeros((2, nb), dtype=torch.float64, device=dev)
    check(_ffi.lib().oevae_hist(ptr(var_t), ptr(w_t), ptr(jzw_t), ptr(score_t), ptr(keep_t), var_t.numel(), float(cut),
                                ptr(edges_t), nb + 1, ptr(counts), ptr(wsums), dev.index, _stream()))
    if reduce:
        _dist.all_reduce_sum(counts); _dist.all_reduce_sum(wsums)
    return counts.cpu().numpy(), wsums.cpu().numpy()


def mass_histograms(sample, plot_var='m', bin_size=2.5, X_loss=None, loss_cut=None, edges=None):
    """Histogram arithmetic of plot_distributions (plots.py:686-715) for one sample dict (host arrays in)."""
    dev = _dev()
    var = np.asarray(sample[plot_var], np.float32)
    if edges is None:
        sel = var if X_loss is None or loss_cut is None else var[np.asarray(X_loss) > loss_cut]
        edges = bin_edges(sel, bin_size)
    sc = None if X_loss is None or loss_cut is None else _to_dev(X_loss, dev)
    c, ws = mass_histograms_device(_to_dev(var, dev), _to_dev(sample['weights'], dev), _to_dev(sample['JZW'], dev),
                                   edges, sc, 0.0 if loss_cut is None else float(loss_cut))
    return edges, c, ws


def generate_cuts(y_true, sample, X_loss, loss_metric='MAE', cut_types=('bkg_eff', 'gain'), plot_vars=('m', 'pt'),
                  bin_sizes={'m': 2.5, 'pt': 10}):
    """plots.py:88-104 without the drawing: for each working point the cut value, the surviving jet count and
    the (uncut, cut) per-class histograms of m and pt.  Returns {name: {...}}."""
    positive_rates = get_rates(y_true, X_loss, sample['weights'])
    points = []
    for cut_type in cut_types:
        if cut_type == 'bkg_eff':
            points += [('bkg_eff_' + format(e, '1.0e'), cut_type, e) for e in [1e-4, 1e-3, 1e-2, 1e-1, 1e0, 1e1]]
        else:
            points += [('best_' + cut_type, cut_type, None)]
    dev = _dev()
    sc = _to_dev(X_loss, dev); w = _to_dev(sample['weights'], dev); jzw = _to_dev(sample['JZW'], dev)
    out = {}
    Xl = np.asarray(X_loss)
    for name, cut_type, eff in points:
        loss_cut = cut_value(y_true, positive_rates, sample['weights'], cut_type, eff)
        mask = Xl > loss_cut
        res = {'loss_cut': float(loss_cut), 'n_pass': int(mask.sum())}
        for pv in plot_vars:
            var = np.asarray(sample[pv], np.float32)
            var_t = _to_dev(var, dev)
            e0 = bin_edges(var, bin_sizes[pv])
            res[pv + '_uncut'] = (e0,) + mass_histograms_device(var_t, w, jzw, e0)
            if mask.any():
                e1 = bin_edges(var[mask], bin_sizes[pv])
                res[pv + '_cut'] = (e1,) + mass_histograms_device(var_t, w, jzw, e1, sc, float(loss_cut))
        out[name] = res
    return out
