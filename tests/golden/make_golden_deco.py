#!/usr/bin/env python
"""Golden fixture for mass_deco (OE-VAE/plots.py:54-85: get_bins, cum_distribution, mass_deco), generated from the REFERENCE's own
function bodies (read from /root/reference with ast and exec'ed in memory; nothing is copied).  Output (committed): ref_deco.npz.
    python tests/golden/make_golden_deco.py"""
import os, sys, textwrap
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import extract, REF


def main():
    from scipy import interpolate
    ns = {'np': np, 'interpolate': interpolate}
    for name, text in extract(REF + '/plots.py', ['get_bins', 'cum_distribution', 'mass_deco']).items():
        exec(compile(textwrap.dedent(text), REF + '/plots.py:' + name, 'exec'), ns)
    rng = np.random.Generator(np.random.Philox(4242))
    N = 40000
    y_true = (rng.uniform(size=N) < 0.8).astype(np.int64)                      # 1 = QCD
    m = np.where(y_true == 1, rng.gamma(2.0, 35.0, N) + 1, rng.normal(172, 15, N)).astype(np.float32).clip(0.5, None)
    pt = (450 + rng.exponential(150, N)).astype(np.float32)
    # a score correlated with the mass (what the decorrelation removes), with ties
    loss = (1 / (1 + np.exp(-(m / 80 - 1 + rng.normal(0, 0.6, N))))).astype(np.float32)
    loss = np.round(loss, 4).astype(np.float32)
    sample = {'m': m, 'pt': pt}
    out = {'y_true': y_true, 'm': m, 'pt': pt, 'loss': loss}
    for deco in ('2d', 'm', 'pt'):
        out['deco_' + deco] = ns['mass_deco'](y_true, sample, loss.copy(), deco=deco)
    out['m_bins'] = ns['get_bins'](m[y_true == 1], True)
    np.savez_compressed(os.path.join(HERE, 'ref_deco.npz'), **out)
    print({k: (v.shape, v.dtype) for k, v in out.items()})


if __name__ == '__main__':
    main()
