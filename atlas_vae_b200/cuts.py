"""Arithmetic of the reference's scoring / cut / histogram chain on the device (no figure drawing).

  score_sample       vae.py:163-167 + plots.py:22-35   (predict x n_iter, loss_function metrics, latent_loss,
                                                        filtering, loss_mapping) in ONE pass over the jets
  loss_mapping       plots.py:15-20
  get_rates          plots.py:228-232   weighted ROC on the device, bit-identical to sklearn.metrics.roc_curve (roc_kernels.cuh)
  best_threshold     plots.py:235-247
  make_cut           plots.py:250-259
  bump_scan_cuts     plots.py:272-282   (the cut grid of bump_scan; the BumpHunter statistics themselves are out of scope)
  mass_histograms    plots.py:697-715   (bin edges from get_idx, per-class weighted histogram, normalize / density scalings)
  mass_deco          plots.py:54-85     (score -> its empirical CDF among the background jets of the same (m, pT) cell)
  generate_cuts      plots.py:88-104    (the six bkg_eff working points + best gain; returns histograms); works on a test set
                                         sharded across ranks: scores are gathered for the threshold search, histogram counts reduced
"""
import ctypes as C

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


# ------------------------------------------------------------------------------------------------ ROC / thresholds
class DeviceROC:
    """roc_curve(y_true, X_loss, pos_label=0, sample_weight=weights) held on the device: float64 fpr / tpr / thresholds exactly as
    sklearn returns them (leading (0, 0, inf) point included)."""

    def __init__(self, fpr, tpr, thr):
        self.fpr, self.tpr, self.thresholds = fpr, tpr, thr

    def rates(self):
        """get_rates' return value (plots.py:230): points with fpr != 0, rates in percent, host float64 arrays."""
        fpr, tpr, thr = (t.cpu().numpy() for t in (self.fpr, self.tpr, self.thresholds))
        sel = fpr != 0
        return 100 * fpr[sel], 100 * tpr[sel], thr[sel]

    def cuts(self, bkg_effs=(), min_tpr=1):
        """Thresholds at the bkg_eff working points (make_cut, plots.py:256-258) and at the best gain (best_threshold, plots.py:237-246),
        found on the device: {'bkg_eff': [loss_cut, ...], 'gain': (loss_cut, gain)}."""
        effs = np.ascontiguousarray(bkg_effs, np.float64)
        out = np.zeros(4 * (len(effs) + 1), np.float64)
        dev = self.fpr.device
        check(_ffi.lib().oevae_roc_cuts(ptr(self.fpr), ptr(self.tpr), ptr(self.thresholds), self.fpr.numel(), ptr(effs) if len(effs) else None,
                                        len(effs), float(min_tpr), ptr(out), dev.index, _stream()))
        out = out.reshape(-1, 4)
        if out[-1, 2] < 0:
            raise ValueError('best_threshold: no ROC point with tpr > %g %%' % min_tpr)     # np.argmax of an empty array in the reference
        return {'bkg_eff': [float(v) for v in out[:-1, 0]], 'gain': (float(out[-1, 0]), float(out[-1, 1]))}

    def searchsorted_cuts(self, eff_val):
        """thresholds[min(searchsorted(fpr, eff_val, 'right'), len - 1)] and the efficiencies there (bump_scan, plots.py:282-283)."""
        q = np.ascontiguousarray(eff_val, np.float64)
        thr, eff = np.zeros(len(q)), np.zeros(len(q))
        check(_ffi.lib().oevae_roc_searchsorted(ptr(self.fpr), ptr(self.thresholds), self.fpr.numel(), ptr(q), len(q), ptr(thr), ptr(eff),
                                                self.fpr.device.index, _stream()))
        return thr, eff


def _gather_ranks(t):
    """Concatenate a per-rank device vector over the ranks in rank order (= the order of the unsharded sample)."""
    import torch.distributed as td
    rank, world = _dist.rank_world()
    if world == 1:
        return t
    n = torch.tensor([t.numel()], dtype=torch.int64, device=t.device)
    ns = [torch.zeros_like(n) for _ in range(world)]
    td.all_gather(ns, n)
    ns = [int(v.item()) for v in ns]
    mx = max(ns)
    pad = torch.zeros(mx, dtype=t.dtype, device=t.device); pad[:t.numel()] = t
    parts = [torch.empty_like(pad) for _ in range(world)]
    td.all_gather(parts, pad)
    return torch.cat([p_[:k] for p_, k in zip(parts, ns)])


def roc_curve_device(y_true, X_loss, weights=None, pos_label=0, gather=True):
    """Weighted ROC on the device.  y_true / X_loss / weights: host arrays or device tensors of this rank's jets; with more than one
    rank the three vectors are gathered first (20M scores = 80 MB), every rank then holds the same curve."""
    dev = _dev() if not isinstance(X_loss, torch.Tensor) else X_loss.device
    s_t = _to_dev(X_loss, dev)
    y_t = (y_true if isinstance(y_true, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(y_true))).to(dev).to(torch.uint8).contiguous()
    w_t = None if weights is None else _to_dev(weights, dev)
    if gather:
        s_t, y_t = _gather_ranks(s_t), _gather_ranks(y_t)
        if w_t is not None: w_t = _gather_ranks(w_t)
    n = s_t.numel()
    fpr = torch.empty(n + 1, dtype=torch.float64, device=dev); tpr = torch.empty_like(fpr); thr = torch.empty_like(fpr)
    npts = C.c_int64(0)
    check(_ffi.lib().oevae_roc(ptr(s_t), ptr(y_t), int(pos_label), ptr(w_t), n, ptr(fpr), ptr(tpr), ptr(thr), C.byref(npts), dev.index, _stream()))
    k = int(npts.value)
    return DeviceROC(fpr[:k], tpr[:k], thr[:k])


def get_rates(y_true, X_loss, weights, metric=None, return_dict=None):
    """plots.py:228-232: (100 fpr, 100 tpr, thresholds) over the points with fpr != 0 -- host arrays, computed on the device."""
    res = roc_curve_device(y_true, X_loss, weights).rates()
    if return_dict is None: return res
    return_dict[metric] = res


def best_threshold(y_true, positive_rates, weights, cut_type, min_tpr=1):
    """plots.py:235-247 on the host arrays get_rates returns."""
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


def bump_scan_cuts(roc, n_cuts=100):
    """The cut grid of bump_scan (plots.py:272-283, eff_type 'bkg'): n_cuts + 1 background efficiencies spaced on a logit scale from
    10^ceil(log10(min fpr)) to 100 %, the thresholds and efficiencies of the ROC points they select."""
    fpr, _, _ = roc.rates()
    logit = lambda x: np.log10(x) - np.log10(1 - x)
    inverse_logit = lambda x: 1 / (1 + 10 ** (-x))
    x_min = 10 ** np.ceil(np.log10(np.min(fpr)))
    eff_val = np.append(100 * inverse_logit(np.linspace(logit(x_min / 100), -logit(x_min / 100), n_cuts)), 100)
    return roc.searchsorted_cuts(eff_val)


# ------------------------------------------------------------------------------------------------ mass decorrelation
def _merge_bins(edges, counts, min_bin_count=2):
    """The merging loop of get_bins (plots.py:58-63) on the occupancy of the 99 initial bins: walking down from the last bin, the first
    bin with fewer than max(2, min_bin_count) jets loses its LEFT edge (it merges into its left neighbour; bin 0 into its right one),
    until the walk reaches bin 0."""
    edges, counts = np.array(edges), [int(c) for c in counts]
    while True:
        idx = 0
        for idx in range(len(edges) - 1)[::-1]:
            if counts[idx] < max(2, min_bin_count):
                edges = np.delete(edges, idx)
                if idx >= 1: counts[idx - 1] += counts[idx]; del counts[idx]
                elif len(counts) > 1: counts[1] += counts[0]; del counts[0]
                else: del counts[0]
                break
        if idx == 0: return edges


def get_bins_device(var_t, keep_t, deco, max_bins=100, min_bin_count=2):
    """get_bins (plots.py:54-63) for the rows of var_t with keep_t: [min, max], or 100 log-spaced edges merged until every bin holds
    at least two of those jets.  Edge arithmetic is NumPy's on the float32 extrema (as the reference's); the device counts."""
    dev = var_t.device
    mm = torch.tensor([np.inf, -np.inf, 0.0], dtype=torch.float64, device=dev)
    check(_ffi.lib().oevae_minmax(ptr(var_t), ptr(keep_t), var_t.numel(), ptr(mm), dev.index, _stream()))
    lo, hi = np.float32(mm[0].item()), np.float32(mm[1].item())
    if not deco: return np.array([lo, hi])
    edges = np.logspace(np.log10(lo), np.log10(hi), num=max_bins)
    e_t = torch.as_tensor(np.asarray(edges, np.float64)).to(dev)
    counts = torch.zeros(len(edges) - 1, dtype=torch.int32, device=dev)
    check(_ffi.lib().oevae_digitize_count(ptr(var_t), ptr(keep_t), var_t.numel(), ptr(e_t), len(edges), ptr(counts), dev.index, _stream()))
    return _merge_bins(edges, counts.cpu().numpy(), min_bin_count)


def mass_deco(y_true, sample, X_loss, deco='2d'):
    """plots.py:68-85: every jet's score becomes its empirical CDF value among the background jets (y_true == 1) of the same (mass, pT)
    cell ('m' / 'pt': one-dimensional binning).  Bin construction on the host from device counts, everything per jet on the device.
    Returns a float32 host array (the reference overwrites X_loss in place and returns it)."""
    dev = _dev()
    m_t, pt_t = _to_dev(sample['m'], dev), _to_dev(sample['pt'], dev)
    loss_t = _to_dev(np.asarray(X_loss, np.float32), dev).clone()
    bkg_t = torch.as_tensor(np.ascontiguousarray(np.asarray(y_true) == 1)).to(dev).to(torch.uint8)
    m_bins = get_bins_device(m_t, bkg_t, deco != 'pt')
    pt_bins = []
    for lo, hi in zip(m_bins[:-1], m_bins[1:]):                                  # get_pt_bins (plots.py:69-70)
        sel = (bkg_t.bool() & (m_t >= float(lo)) & (m_t < float(hi))).to(torch.uint8)
        pt_bins.append(get_bins_device(pt_t, sel, deco != 'm'))
    off = np.zeros(len(m_bins), np.int32); off[1:] = np.cumsum([len(b) for b in pt_bins])
    me_t = torch.as_tensor(np.asarray(m_bins, np.float64)).to(dev)
    pe_t = torch.as_tensor(np.concatenate(pt_bins).astype(np.float64)).to(dev)
    off_t = torch.as_tensor(off).to(dev)
    check(_ffi.lib().oevae_mass_deco(ptr(loss_t), ptr(m_t), ptr(pt_t), ptr(bkg_t), loss_t.numel(), ptr(me_t), len(m_bins), ptr(pe_t), ptr(off_t),
                                     dev.index, _stream()))
    return _host(loss_t)


# ------------------------------------------------------------------------------------------------ histograms
def bin_edges(var, bin_size):
    """plots.py:697-700: get_idx(max, bin_size, min_val=max(0,min), integer=False, tuples=False) -> float64 edges."""
    return get_idx(np.max(var), bin_size=bin_size, min_val=max(0, np.min(var)), integer=False, tuples=False)


def _minmax_device(var_t, score_t=None, cut=None, keep_t=None):
    """(min, max) of var over the rows with keep and score > cut, all-reduced over ranks; float32 values like np.min / np.max."""
    sel = torch.ones(var_t.numel(), dtype=torch.bool, device=var_t.device)
    if keep_t is not None: sel &= keep_t.bool()
    if score_t is not None and cut is not None: sel &= score_t > cut
    v = var_t[sel]
    lo = v.min().reshape(1).double() if v.numel() else torch.full((1,), np.inf, dtype=torch.float64, device=var_t.device)
    hi = v.max().reshape(1).double() if v.numel() else torch.full((1,), -np.inf, dtype=torch.float64, device=var_t.device)
    _dist.all_reduce_minmax(lo, hi)
    return np.float32(lo.item()), np.float32(hi.item())


def total_weight(w_t, keep_t=None):
    """np.sum(sample['weights']) of plots.py:707 as a float32 scalar: float64 device sum (all-reduced), rounded once."""
    acc = torch.zeros(1, dtype=torch.float64, device=w_t.device)
    check(_ffi.lib().oevae_wsum(ptr(w_t), ptr(keep_t), w_t.numel(), ptr(acc), w_t.device.index, _stream()))
    _dist.all_reduce_sum(acc)
    return np.float32(acc.item())


def mass_histograms_device(var_t, w_t, jzw_t, edges, score_t=None, cut=0.0, keep_t=None, reduce=True, norm_weight=None, density=False):
    """K5: per-class (row 0: JZW==-1, row 1: JZW>=0) histogram of `var` for the jets with score > cut.
    norm_weight: total weight of the UNCUT sample (plots.py:707, weight_type 'None') -> weights are scaled by fp32(100) / norm_weight;
    density: weights divided by the bin width (plots.py:709-712).  Returns (counts int64 [2,nb], wsums float64 [2,nb]) on the host;
    summed over ranks when `reduce`."""
    dev = var_t.device
    edges_t = torch.as_tensor(np.ascontiguousarray(edges, np.float64)).to(dev)
    nb = edges_t.numel() - 1
    if nb < 1:                                  # a selection with min == max gives a single edge (utils.py:541-548): no bins
        return np.zeros((2, 0), np.int64), np.zeros((2, 0), np.float64)
    counts = torch.zeros((2, nb), dtype=torch.int64, device=dev)
    wsums = torch.zeros((2, nb), dtype=torch.float64, device=dev)
    wscale = float(np.float32(100) / np.float32(norm_weight)) if norm_weight is not None else 1.0
    check(_ffi.lib().oevae_hist(ptr(var_t), ptr(w_t), ptr(jzw_t), ptr(score_t), ptr(keep_t), var_t.numel(), float(cut),
                                ptr(edges_t), nb + 1, wscale, 1 if density else 0, ptr(counts), ptr(wsums), dev.index, _stream()))
    if reduce:
        _dist.all_reduce_sum(counts); _dist.all_reduce_sum(wsums)
    return counts.cpu().numpy(), wsums.cpu().numpy()


def mass_histograms(sample, plot_var='m', bin_size=2.5, X_loss=None, loss_cut=None, edges=None, normalize=False, density=False):
    """Histogram arithmetic of plot_distributions (plots.py:686-715) for one sample dict (host arrays in).  normalize / density:
    the `weights *= 100 / sum(weights)` and `weights /= bin width` scalings of plots.py:706-712 (both True in the reference's call)."""
    dev = _dev()
    var = np.asarray(sample[plot_var], np.float32)
    if edges is None:
        sel = var if X_loss is None or loss_cut is None else var[np.asarray(X_loss) > loss_cut]
        edges = bin_edges(sel, bin_size)
    sc = None if X_loss is None or loss_cut is None else _to_dev(X_loss, dev)
    w_t = _to_dev(sample['weights'], dev)
    c, ws = mass_histograms_device(_to_dev(var, dev), w_t, _to_dev(sample['JZW'], dev),
                                   edges, sc, 0.0 if loss_cut is None else float(loss_cut),
                                   norm_weight=total_weight(w_t) if normalize else None, density=density)
    return edges, c, ws


def generate_cuts(y_true, sample, X_loss, loss_metric='MAE', cut_types=('bkg_eff', 'gain'), plot_vars=('m', 'pt'),
                  bin_sizes={'m': 2.5, 'pt': 10}):
    """plots.py:88-104 without the drawing: for each working point the cut value, the surviving jet count and the (uncut, cut)
    per-class histograms of m and pt -- integer counts, and the weighted / normalised / density values the reference plots
    (plots.py:706-715).  `sample`, `y_true`, `X_loss` are THIS RANK's shard: the threshold search runs on the gathered scores, every
    rank cuts its own jets, counts and weight sums are SUM-reduced (bit-exact for the integer counts).
    Returns {name: {'loss_cut', 'n_pass', '<var>_uncut': (edges, counts, density), '<var>_cut': ...}}."""
    dev = _dev()
    sc = _to_dev(X_loss, dev); w = _to_dev(sample['weights'], dev); jzw = _to_dev(sample['JZW'], dev)
    roc = roc_curve_device(y_true, sc, w)
    effs = [1e-4, 1e-3, 1e-2, 1e-1, 1e0, 1e1] if 'bkg_eff' in cut_types else []
    cuts = roc.cuts(effs)
    points = [('bkg_eff_' + format(e, '1.0e'), c) for e, c in zip(effs, cuts['bkg_eff'])]
    if 'gain' in cut_types: points.append(('best_gain', cuts['gain'][0]))
    norm = total_weight(w)                                              # plots.py:707: the uncut sample's total weight (weight_type 'None')
    var_t = {pv: _to_dev(np.asarray(sample[pv], np.float32), dev) for pv in plot_vars}
    uncut = {}
    for pv in plot_vars:
        lo, hi = _minmax_device(var_t[pv])
        e0 = get_idx(hi, bin_size=bin_sizes[pv], min_val=max(0, lo), integer=False, tuples=False)
        uncut[pv] = (e0,) + mass_histograms_device(var_t[pv], w, jzw, e0, norm_weight=norm, density=True)
    out = {}
    for name, loss_cut in points:
        n_pass = (sc > loss_cut).sum().reshape(1); _dist.all_reduce_sum(n_pass)
        res = {'loss_cut': float(loss_cut), 'n_pass': int(n_pass.item())}
        for pv in plot_vars:
            res[pv + '_uncut'] = uncut[pv]
            if res['n_pass'] > 0:
                # plots.py:697: the bins of the cut sample span ITS OWN range (sample = samples[1], all classes); the weights are
                # still normalised by the uncut sample's total (plots.py:707: samples[0])
                lo, hi = _minmax_device(var_t[pv], sc, float(loss_cut))
                e1 = get_idx(hi, bin_size=bin_sizes[pv], min_val=max(0, lo), integer=False, tuples=False)
                res[pv + '_cut'] = (e1,) + mass_histograms_device(var_t[pv], w, jzw, e1, sc, float(loss_cut), norm_weight=norm, density=True)
        out[name] = res
    return out
