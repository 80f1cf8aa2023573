#!/usr/bin/env python
"""Golden fixture for the OoD <-> background pairing (OE-VAE/utils.py:193-227) and reweighting (utils.py:230-274), generated from the
REFERENCE's own function bodies (read from /root/reference with ast and exec'ed in memory, as make_golden.py does; nothing is copied).

np.random.choice is replaced by a recorder for the duration of the call, so that the fixture holds, for every background jet, the
candidate SET the reference would draw from (first / last candidate, count) and the window widths it ended with -- the draw itself
(np.random under fork()) is not reproducible.  Output (committed): ref_pairing.npz.      python tests/golden/make_golden_pairing.py
"""
import os, sys, textwrap, time
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import extract, REF


def main():
    import multiprocessing as mp
    ns = {'np': np, 'time': time, 'mp': mp}
    for name, text in extract(REF + '/utils.py', ['OoD_pairing', 'reweight_sample', 'get_weights', 'get_idx']).items():
        exec(compile(textwrap.dedent(text), REF + '/utils.py:' + name, 'exec'), ns)
    rng = np.random.Generator(np.random.Philox(777))
    n_bkg, n_ood = 3000, 4000
    bkg = {'m': rng.gamma(2.0, 35.0, n_bkg).astype(np.float32), 'pt': rng.uniform(450, 1200, n_bkg).astype(np.float32),
           'weights': rng.uniform(0.5, 1.5, n_bkg).astype(np.float32)}
    ood = {'m': rng.normal(80.4, 9.0, n_ood).astype(np.float32), 'pt': (450 + rng.exponential(120, n_ood)).astype(np.float32),
           'weights': rng.uniform(0.5, 1.5, n_ood).astype(np.float32), 'id': np.arange(n_ood)}
    bkg['m'][:5] = [0.0, 900.0, 80.4, 79.9, 300.0]; bkg['pt'][:5] = [450.0, 3000.0, 500.0, 2500.0, 460.0]      # far-away jets: many widenings
    rec = []
    real_choice = np.random.choice
    def recorder(a, *args, **kw):
        a = np.asarray(a); rec.append((int(a[0]), int(a[-1]), int(a.size))); return a[0]
    np.random.choice = recorder
    try:
        paired = ns['OoD_pairing'](dict(bkg), dict(ood), multithread=False, verbose=False, seed=0)
    finally:
        np.random.choice = real_choice
    rec = np.array(rec, np.int64)
    order = np.argsort(ood['m'])
    # the recorder returns the first candidate: paired['id'] = id of the first candidate in the m-sorted sample
    out = {'bkg_m': bkg['m'], 'bkg_pt': bkg['pt'], 'bkg_w': bkg['weights'], 'ood_m': ood['m'], 'ood_pt': ood['pt'], 'ood_w': ood['weights'],
           'order': order, 'first_rel': rec[:, 0], 'last_rel': rec[:, 1], 'count': rec[:, 2], 'paired_first_id': paired['id']}
    # reweight_sample 'X-S' (utils.py:248-249) and 'none' (utils.py:240-242)
    b2, s2 = ns['reweight_sample']({k: v.copy() for k, v in bkg.items()}, {k: v.copy() for k, v in ood.items()}, {'m': 10, 'pt': 50}, 'X-S')
    out['xs_sig_w'] = s2['weights']
    for wt in ('flat_2d', 'OoD_2d', 'flat_m', 'OoD_pt'):
        b3, s3 = ns['reweight_sample']({k: v.copy() for k, v in bkg.items()}, {k: v.copy() for k, v in ood.items()}, {'m': 10, 'pt': 50}, wt)
        out['rw_%s_sig' % wt] = s3['weights']; out['rw_%s_bkg' % wt] = b3['weights']
    np.savez_compressed(os.path.join(HERE, 'ref_pairing.npz'), **out)
    print('ref_pairing.npz:', {k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    main()
