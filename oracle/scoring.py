"""NumPy restatement of the anomaly-scoring / cut / histogram chain (TEST INFRASTRUCTURE, see oracle/__init__.py).

  predict (+ n_iter mean)   vae.py:163-166        -> predict()
  loss_function             utils.py:346-375      -> score_metrics()
  latent_loss               utils.py:378-386      -> latent_score()
  filtering                 utils.py:532-538      -> finite_rows()
  loss_mapping              plots.py:15-20        -> loss_mapping()
  get_rates                 plots.py:228-232      -> get_rates()            (real sklearn roc_curve)
  best_threshold            plots.py:235-247      -> best_threshold()
  make_cut                  plots.py:250-259      -> cut_threshold() / cut_mask()
  get_idx (bin edges)       utils.py:541-548      -> bin_edges()
  plot_distributions hist   plots.py:697-715      -> class_histograms()     (real np.histogram)

PINNED against the reference's own function bodies (tests/golden/make_golden.py -> tests/golden/ref_*.npz).
"""
import warnings
import numpy as np

from . import vae_numpy as V


def predict(flat, FC_layers, input_dim, X, eps_list, dtype=np.float32, batch=10000):
    """model.predict averaged over n_iter noise draws (vae.py:163-166): mean of X_pred BEFORE metrics."""
    n_hidden = len(FC_layers) - 1
    layers = [(W.astype(dtype), b.astype(dtype)) for W, b in V.unflatten(np.asarray(flat), FC_layers, input_dim)]
    X = np.asarray(X, dtype)
    out = np.zeros((len(eps_list),) + X.shape, np.float32)
    kld = None
    for n, eps in enumerate(eps_list):
        for i in range(0, len(X), batch):
            c = V.forward(layers, X[i:i + batch], np.asarray(eps[i:i + batch], dtype), n_hidden)
            out[n, i:i + batch] = c['y']
    return np.mean(out, axis=0, dtype=np.float32) if len(eps_list) > 1 else out[0]


def latent_score(flat, FC_layers, input_dim, X, dtype=np.float32):
    """latent_loss (utils.py:378-386): per-jet latent KLD of a forward pass, non-finite -> 0."""
    n_hidden = len(FC_layers) - 1
    layers = [(W.astype(dtype), b.astype(dtype)) for W, b in V.unflatten(np.asarray(flat), FC_layers, input_dim)]
    c = V.forward(layers, np.asarray(X, dtype), None, n_hidden, encoder_only=True)
    k = c['kld']
    return np.where(np.isfinite(k), k, 0)


def _kld_terms(P, Q):
    with warnings.catch_warnings(), np.errstate(all='ignore'):
        warnings.simplefilter('ignore')
        return np.nan_to_num(P * np.log2(P / Q))            # NaN->0, +-inf -> +-finfo.max  (utils.py:347-350)


def score_metrics(P, Q, metrics=('MSE', 'MAE', 'KLD', 'JSD')):
    """loss_function (utils.py:346-375) for the metrics on the path; computed in the dtype of P/Q."""
    out = {}
    with np.errstate(all='ignore'):
        for m in metrics:
            if m == 'MSE':
                out[m] = np.mean((P - Q) ** 2, axis=1)
            elif m == 'MAE':
                out[m] = np.mean(np.abs(P - Q), axis=1)
            elif m == 'MARE':
                out[m] = np.mean(np.abs(P - Q) / P, axis=1)
            elif m == 'KLD':
                out[m] = np.sum(_kld_terms(P, Q), axis=1)
            elif m == 'JSD':
                M = (P + Q) / 2
                out[m] = np.sum((_kld_terms(P, M) + _kld_terms(Q, M)) / 2, axis=1)
            elif m == 'X-S':
                out[m] = np.sum(_kld_terms(P, P * Q), axis=1)
            else:
                raise ValueError(m)
    return out


def finite_rows(X_pred):
    """filtering (utils.py:532-538): True for rows to KEEP (all of X_pred finite)."""
    return np.all(np.isfinite(X_pred), axis=1)


def loss_mapping_branch(x):
    """Which of the five global branches of loss_mapping (plots.py:15-20) applies."""
    if np.all(np.logical_and(x >= 0, x <= 1)): return 0
    if np.all(np.logical_and(x >= -1, x <= 0)): return 1
    if np.all(x >= 0): return 2
    if np.all(x <= 0): return 3
    return 4


def loss_mapping(x):
    b = loss_mapping_branch(x)
    if b == 0: return x
    if b == 1: return x + 1
    if b == 2: return x / (np.abs(x) + 1)
    if b == 3: return x / (np.abs(x) + 1) + 1
    return (x / (np.abs(x) + 1) + 1) / 2


def get_rates(y_true, X_loss, weights):
    from sklearn import metrics
    fpr, tpr, thr = metrics.roc_curve(y_true, X_loss, pos_label=0, sample_weight=weights)
    keep = fpr != 0
    return 100 * fpr[keep], 100 * tpr[keep], thr[keep]


def best_threshold(positive_rates, cut_type='gain', min_tpr=1, y_true=None, weights=None):
    fpr, tpr, thr = positive_rates
    sel = tpr > min_tpr
    fpr, tpr, thr = fpr[sel], tpr[sel], thr[sel]
    if cut_type == 'gain':
        vals, factor = tpr / fpr, 1
    else:
        vals = tpr / np.sqrt(fpr)
        factor = np.sum(weights[y_true == 0]) / np.sqrt(np.sum(weights[y_true == 1])) / 10
    i = np.argmax(vals)
    return thr[i], vals[i] * factor


def cut_threshold(positive_rates, bkg_eff=None, cut_type='gain', y_true=None, weights=None):
    """Threshold part of make_cut (plots.py:250-258)."""
    if bkg_eff is None:
        return best_threshold(positive_rates, cut_type, y_true=y_true, weights=weights)[0]
    fpr, tpr, thr = positive_rates
    return thr[np.argmin(abs(fpr - bkg_eff))]


def cut_mask(X_loss, loss_cut):
    """Selection of make_cut (plots.py:259): strictly greater; NaN scores fail the cut."""
    with np.errstate(invalid='ignore'):
        return X_loss > loss_cut


def bin_edges(var_all, bin_size):
    """get_idx(max, bin_size, min_val=max(0,min), integer=False, tuples=False)  (utils.py:541-548, plots.py:697-700):
    float64 edges np.append(np.arange(min,max,step),max); the last bin is ragged."""
    min_val, max_val = max(0, np.min(var_all)), np.max(var_all)
    return np.append(np.arange(min_val, max_val, bin_size), max_val)


def class_histograms(var, weights, JZW, edges, mask=None):
    """Per-class counts and weighted sums of plot_distributions (plots.py:686-715), arithmetic only.
    class 0: JZW == -1 (signal), class 1: JZW >= 0 (QCD).  Returns counts int64 [2,nb], wsums float64 [2,nb]
    of the raw per-jet weights (the 100/sum(w)/binwidth normalisation is a per-bin host rescale)."""
    nb = len(edges) - 1
    counts = np.zeros((2, nb), np.int64); wsums = np.zeros((2, nb), np.float64)
    keep = np.ones(len(var), bool) if mask is None else np.asarray(mask, bool)
    for c, cond in enumerate([JZW == -1, JZW >= 0]):
        sel = cond & keep
        v = np.float32(var[sel])
        counts[c] = np.histogram(v, edges)[0]
        wsums[c] = np.histogram(v, edges, weights=np.float64(weights[sel]))[0]
    return counts, wsums
