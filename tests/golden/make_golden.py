#!/usr/bin/env python
"""Generate the golden fixtures in tests/golden/ from the REFERENCE's own code.

Run in the build container (where /root/reference exists):   python tests/golden/make_golden.py

The reference modules import tensorflow / h5py / matplotlib / energyflow at the top, none of which is
installed, so they cannot be imported.  Their pure NumPy / sklearn functions CAN run: this script pulls the
function bodies out of the reference sources with ``ast`` (nothing is copied into the repo -- the text is
read from /root/reference at generation time and exec'ed in memory) and evaluates them on small seeded inputs.
Outputs (committed):  ref_scoring.npz, ref_preprocess.npz, ref_checkpoint.json.
The TensorFlow-resident arithmetic (models.py:7-132,181-186) cannot be executed this way: the second half
of this script therefore writes SELF-generated vectors (oracle float64) to oracle_train.npz, which pin the
oracle against regressions only ("parity unpinned" for that part).
"""
import ast, io, json, os, sys, contextlib, warnings
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference/OE-VAE'
sys.path.insert(0, os.path.join(HERE, '..', '..'))


def extract(path, names):
    """source text of the (possibly nested) function defs ``names`` in ``path``."""
    src = open(path).read()
    tree = ast.parse(src)
    found = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in names and node.name not in found:
            found[node.name] = ast.get_source_segment(src, node)
    missing = set(names) - set(found)
    assert not missing, missing
    return found


def load_ref():
    from sklearn import metrics
    ns = {'np': np, 'warnings': warnings, 'metrics': metrics}
    import textwrap
    for path, names in [(REF + '/utils.py', ['loss_function', 'get_idx', 'jets_sorting', 'get_4v', 'filtering']),
                        (REF + '/plots.py', ['loss_mapping', 'get_rates', 'best_threshold', 'make_cut']),
                        (REF + '/models.py', ['model_checkpoint'])]:
        for name, text in extract(path, names).items():
            exec(compile(textwrap.dedent(text), path + ':' + name, 'exec'), ns)
    return ns


def main():
    ref = load_ref()
    rng = np.random.Generator(np.random.Philox(20211213))

    # ---------------- scoring metrics (utils.py:346-375), filtering (:532-538), loss_mapping (plots.py:15-20)
    N, D = 64, 24
    P = rng.normal(0.2, 1.0, (N, D)).astype(np.float32)
    Q = (P + rng.normal(0, 0.3, (N, D))).astype(np.float32)
    P[0, :4] = 0; Q[1, :4] = 0; P[2, 3] = Q[2, 3]; Q[3, 5] = -P[3, 5]        # 0/0, x/0, equal, sign flips
    P[4] = np.abs(P[4]); Q[4] = np.abs(Q[4])                                   # an all-positive row
    out = {'P': P, 'Q': Q}
    for metric in ['MSE', 'MAE', 'MARE', 'KLD', 'JSD', 'X-S']:
        with np.errstate(all='ignore'):
            out['m_' + metric] = ref['loss_function'](P, Q, 4, metric, multiloss=False)
    Qbad = Q.copy(); Qbad[5, 2] = np.nan; Qbad[9, 0] = np.inf; Qbad[9, 7] = -np.inf
    y = np.arange(N); samp = {'m': np.arange(N, dtype=np.float32)}
    y2, Xt2, Xp2, s2 = ref['filtering'](y, P.copy(), Qbad.copy(), samp)
    out['filt_bad'] = Qbad; out['filt_kept'] = y2
    for i, v in enumerate([rng.uniform(0, 1, 50), rng.uniform(-1, 0, 50), rng.uniform(0, 30, 50),
                           rng.uniform(-30, 0, 50), rng.normal(0, 5, 50)]):
        v = v.astype(np.float32)
        out['map_in%d' % i] = v; out['map_out%d' % i] = ref['loss_mapping'](v)

    # ---------------- ROC / thresholds / cut (plots.py:228-259)
    M = 4000
    y_true = (rng.uniform(size=M) < 0.7).astype(np.int64)                      # 1 = QCD, 0 = signal (vae.py:150)
    score = np.where(y_true == 1, rng.beta(2, 6, M), rng.beta(4, 3, M)).astype(np.float32)
    score[::97] = score[1::97][:len(score[::97])]                              # ties
    wts = np.where(y_true == 1, 1.0, 1e-3).astype(np.float32) * rng.uniform(0.5, 1.5, M).astype(np.float32)
    fpr, tpr, thr = ref['get_rates'](y_true, score, wts)
    out.update(roc_y=y_true, roc_score=score, roc_w=wts, roc_fpr=fpr, roc_tpr=tpr, roc_thr=thr)
    sample = {'weights': wts, 'idx': np.arange(M)}
    with contextlib.redirect_stdout(io.StringIO()):
        thr_gain, val_gain = ref['best_threshold'](y_true, (fpr, tpr, thr), wts, 'gain')
        kept_gain = ref['make_cut'](y_true, score, sample, (fpr, tpr, thr), 'MAE', 'gain')['idx']
    out.update(thr_gain=thr_gain, val_gain=val_gain, kept_gain=kept_gain)
    for eff in [1e-1, 1e0, 1e1]:
        kept = ref['make_cut'](y_true, score, sample, (fpr, tpr, thr), 'MAE', 'bkg_eff', eff)['idx']
        out['kept_eff_%g' % eff] = kept

    # ---------------- bin edges (utils.py:541-548 as called at plots.py:697-700) + histogram (np.histogram under pylab.hist)
    mass = np.concatenate([rng.gamma(3.0, 30.0, M - 4), [0.0, 2.5, 5.0, 617.25]]).astype(np.float32)
    mass[10] = np.float32(mass.max())
    for bs, tag in [(2.5, 'm'), (10, 'pt')]:
        var = mass if tag == 'm' else (mass * np.float32(3.7) + np.float32(450)).astype(np.float32)
        edges = ref['get_idx'](np.max(var), bin_size=bs, min_val=max(0, np.min(var)), integer=False, tuples=False)
        out['var_' + tag] = var; out['edges_' + tag] = edges
        out['hist_' + tag] = np.histogram(var, edges)[0]
    out['idx_tuples'] = np.array(ref['get_idx'](1003, 7))
    np.savez_compressed(os.path.join(HERE, 'ref_scoring.npz'), **out)

    # ---------------- preprocessing (utils.py:564-571, 300-308)
    Nj, n_raw = 96, 40
    raw = np.zeros((Nj, n_raw, 4), np.float32)
    for i in range(Nj):
        n = int(rng.integers(1, n_raw + 1))
        p = rng.normal(0, 20, (n, 3)).astype(np.float32)
        raw[i, :n, 1:] = p
        raw[i, :n, 0] = np.sqrt((p.astype(np.float64) ** 2).sum(1)).astype(np.float32) * np.float32(1.0001)
        raw[i, :n] = raw[i, rng.permutation(n)]
    raw = raw.reshape(Nj, -1)
    srt = ref['jets_sorting'](raw)
    fv = ref['get_4v'](srt[:, :4 * 20])
    np.savez_compressed(os.path.join(HERE, 'ref_preprocess.npz'), raw=raw, sorted=srt, pt20=fv['pt_calo'], m20=fv['m_calo'],
                        raw16=raw.astype(np.float16), sorted16=ref['jets_sorting'](raw.astype(np.float16)))

    # ---------------- model_checkpoint state machine (models.py:222-240)
    class Var:
        def __init__(s, v): s.v = v
        def numpy(s): return s.v
        def __lt__(s, o): return s.v < o
        def __truediv__(s, o): return Var(s.v / o)
    class Opt: pass
    class Vae:
        def __init__(s): s.saved = 0
        def save_weights(s, path): s.saved += 1
    cases = []
    for series in [[5, 4, 3.5, 3.4995, 3.4999, 3.4998, 3.4997, 3.4996, 3.4996, 3.4996, 3.4995, 3.4995],
                   [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18],
                   [3, 2.9, 2.95, 2.7, 2.71, 2.72, 2.73, 2.5, 2.6, 2.6, 2.6, 2.6]]:
        opt = Opt(); opt.learning_rate = Var(1e-3 if series[0] != 1 else 1.5e-4); vae = Vae(); count = 0; trace = []
        for ep in range(1, len(series)):
            hist = {'Train loss': series[:ep + 1]}
            with contextlib.redirect_stdout(io.StringIO()):
                opt, count = ref['model_checkpoint'](vae, opt, hist, 'x.h5', count)
            trace.append([vae.saved, opt.learning_rate.v, -1 if count is None else count])
            if count is None: break
        cases.append({'series': series, 'lr0': 1e-3 if series[0] != 1 else 1.5e-4, 'trace': trace})
    json.dump(cases, open(os.path.join(HERE, 'ref_checkpoint.json'), 'w'), indent=1)

    # ---------------- self-generated training vectors (oracle float64; regression pin only)
    from oracle import vae_numpy as V, philox
    FC, Din, B = [80, 40, 10], 80, 256
    flat = V.init_params(FC, Din, seed=0)
    X = rng.normal(0, 0.5, (B, Din)).astype(np.float32); X[:, 60:] = 0
    Xo = rng.normal(0.1, 0.7, (B, Din)).astype(np.float32)
    w = rng.uniform(0.5, 1.5, B).astype(np.float32); wo = rng.uniform(0.5, 1.5, B).astype(np.float32)
    eps = philox.normal_eps(0, B, FC[-1])
    tr = {'FC': np.array(FC), 'D': Din, 'flat0': flat, 'X': X, 'Xo': Xo, 'w': w, 'wo': wo, 'eps': eps}
    for oe in V.OE_TYPES:
        (r, k, o, t), g = V.losses_and_grad(flat, FC, Din, X, w, Xo, wo, eps, OE_type=oe, beta=0.5, lamb=2.0, margin=0.3)
        tr['loss_' + oe] = np.stack([r, k, o, t]); tr['grad_' + oe] = g
    T = V.Trainer(flat, FC, Din, 'KLD')
    curve = []
    for s in range(20):
        o4 = T.step(X, w, Xo, wo, eps, lr=1e-3, beta=0.5, lamb=2.0, margin=0.3)
        curve.append([float(np.mean(q)) for q in o4])
        if s == 0: tr['theta1'] = T.theta.copy()
    tr['theta20'] = T.theta.copy(); tr['curve20'] = np.array(curve)
    np.savez_compressed(os.path.join(HERE, 'oracle_train.npz'), **tr)
    print('golden fixtures written to', HERE)


if __name__ == '__main__':
    main()
