"""CPU tests of the host-side logic that mirrors the reference (no GPU needed)."""
import json, os
import numpy as np


def test_model_checkpoint_matches_reference(golden_dir):
    from atlas_vae_b200.models import model_checkpoint, Adam

    class FakeVae:
        def __init__(s): s.saved = 0
        def save_weights(s, p): s.saved += 1
    for case in json.load(open(os.path.join(golden_dir, 'ref_checkpoint.json'))):
        opt, vae, count = Adam(case['lr0']), FakeVae(), 0
        sink = open(os.devnull, 'w')
        for ep, exp in zip(range(1, len(case['series'])), case['trace']):
            opt, count = model_checkpoint(vae, opt, {'Train loss': case['series'][:ep + 1]}, 'x.h5', count, _out=sink)
            assert [vae.saved, opt.learning_rate, -1 if count is None else count] == exp
            if count is None: break


def test_init_params_and_layout_match_oracle():
    from atlas_vae_b200.models import init_params
    from oracle import vae_numpy as V
    for FC, D in [([80, 40, 20, 10], 400), ([16], 12), ([1024, 512, 256, 64], 400)]:
        a = init_params(FC, D, 0)
        assert a.dtype == np.float32 and np.array_equal(a, V.init_params(FC, D, 0))


def test_cut_host_logic_matches_reference(golden_dir):
    from atlas_vae_b200 import cuts as Cc, utils as U
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    # the curve itself is computed on the device (tests/test_gpu_roc.py pins it to the same golden arrays); here the host half of
    # make_cut / best_threshold runs on the reference's curve
    rates = (g['roc_fpr'], g['roc_tpr'], g['roc_thr'])
    thr, val = Cc.best_threshold(g['roc_y'], rates, g['roc_w'], 'gain')
    assert thr == g['thr_gain'] and val == g['val_gain']
    samp = {'weights': g['roc_w'], 'idx': np.arange(len(g['roc_y']))}
    np.testing.assert_array_equal(Cc.make_cut(g['roc_y'], g['roc_score'], samp, rates, 'MAE', 'gain')['idx'], g['kept_gain'])
    for tag, bs in [('m', 2.5), ('pt', 10)]:
        np.testing.assert_array_equal(Cc.bin_edges(g['var_' + tag], bs), g['edges_' + tag])
    np.testing.assert_array_equal(np.array(U.get_idx(1003, 7)), g['idx_tuples'])
    for i, (lo, hi) in enumerate([(0.1, 0.9), (-0.8, -0.1), (0.0, 30.), (-30., 0.), (-5., 5.)]):
        assert Cc.loss_mapping_branch(lo, hi) == i
    assert Cc.loss_mapping_branch(0.1, 0.9, True) == 4


def test_filtering_matches_reference(golden_dir):
    from atlas_vae_b200 import utils as U
    g = np.load(os.path.join(golden_dir, 'ref_scoring.npz'))
    y, Xt, Xp, s = U.filtering(np.arange(len(g['P'])), g['P'].copy(), g['filt_bad'].copy(), {'m': np.arange(len(g['P']))})
    np.testing.assert_array_equal(y, g['filt_kept'])
    assert len(Xt) == len(Xp) == len(s['m']) == len(y)


def test_synthetic_jets_are_deterministic():
    from atlas_vae_b200 import synthetic as S
    a, ja = S.make_jets('W', 300, 60, seed=3)
    b, jb = S.make_jets('W', 300, 60, seed=3)
    assert a.shape == (300, 240) and bool((a == b).all()) and float(ja[0]) == -1.0
    v = a.view(300, 60, 4)
    assert bool((v[:, :, 0] >= 0).all())
    m2 = v[:, :, 0] ** 2 - (v[:, :, 1:] ** 2).sum(-1)
    assert float(m2.abs().max()) < 1e-1 * float((v[:, :, 0] ** 2).max())          # massless constituents


def test_cli_accepts_reference_flags_and_readme_aliases():
    """vae.py:16-66 flag names and the README's older spellings land on the same fields."""
    from atlas_vae_b200.vae import parse_args
    a = parse_args(['--n_train', '1e5', '--n_const', '20', '--n_dims', '4', '--FC_layers', '80', '40', '10', '--batch_size', '5e3',
                    '--beta', '1', '--apply_cuts', 'ON', '--output_dir', 'o'])
    assert (a.n_train, a.batch_size, a.FC_layers, a.apply_cuts) == (100000, 5000, [80, 40, 10], 'ON')
    assert a.model_out == 'o/model.h5' and a.hist_file == 'o/history.pkl'
    b = parse_args(['--n_constituents', '20', '--FCN_neurons', '80', '40', '--latent_dim', '10', '--apply_cut', 'ON',
                    '--n_test', '2e5', '--n_W', '3e5', '--n_top', '4e5', '--patience', '5', '--n_gpus', '8'])
    assert b.n_const == 20 and b.FC_layers == [80, 40, 10] and b.apply_cuts == 'ON'
    assert (b.n_valid, b.n_OoD, b.n_sig, b.patience, b.n_gpus) == (200000, 300000, 400000, 5, 8)
    assert parse_args([]).FC_layers == [80, 40, 20, 10] and parse_args(['--latent_dim', '6']).FC_layers == [80, 40, 20, 6]


def test_reweight_sample_matches_reference(golden_dir):
    """reweight_sample / get_weights (utils.py:230-274) against outputs of the reference's own function bodies
    (tests/golden/make_golden_pairing.py); host NumPy, needs no GPU."""
    import importlib.util, sys, types
    g = np.load(os.path.join(golden_dir, 'ref_pairing.npz'))
    # atlas_vae_b200.utils imports torch-free helpers lazily; load the two functions without touching the GPU
    from atlas_vae_b200 import utils as U
    mk_b = lambda: {'m': g['bkg_m'].copy(), 'pt': g['bkg_pt'].copy(), 'weights': g['bkg_w'].copy()}
    mk_s = lambda: {'m': g['ood_m'].copy(), 'pt': g['ood_pt'].copy(), 'weights': g['ood_w'].copy()}
    _, s = U.reweight_sample(mk_b(), mk_s(), {'m': 10, 'pt': 50}, 'X-S')
    np.testing.assert_array_equal(s['weights'], g['xs_sig_w'])
    for wt in ('flat_2d', 'OoD_2d', 'flat_m', 'OoD_pt'):
        b, s = U.reweight_sample(mk_b(), mk_s(), {'m': 10, 'pt': 50}, wt)
        np.testing.assert_array_equal(s['weights'], g['rw_%s_sig' % wt]); np.testing.assert_array_equal(b['weights'], g['rw_%s_bkg' % wt])
    b, s = U.reweight_sample(mk_b(), mk_s(), {'m': 10, 'pt': 50}, 'None')
    assert np.all(b['weights'] == 1) and np.all(s['weights'] == 1)
