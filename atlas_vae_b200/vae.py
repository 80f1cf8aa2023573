"""vae.py-compatible driver (OE-VAE/vae.py:15-176) on the B200 path:  python -m atlas_vae_b200.vae [flags]

Same flags, defaults and path handling as the reference (vae.py:16-66), plus the README's older spellings as
aliases (README.md:51-67): --n_constituents, --FCN_neurons + --latent_dim, --apply_cut, --n_test, --n_W,
--n_top, --patience; new: --n_gpus (informational: launch with torchrun, one process per GPU), --precision, --seed.
There is no HDF5 data in this environment (h5py absent), so samples are the deterministic synthetic jets of
SURVEY 8(d) (QCD background, W out-of-distribution sample, top signal); the control flow after loading follows
vae.py:89-176 without the figures: scaler fit/apply, train_model, scoring pass, cuts and jet-mass histograms.
"""
import os
import pickle
import sys
from argparse import ArgumentParser
from pathlib import Path

import numpy as np


def parse_args(argv=None):
    p = ArgumentParser()
    p.add_argument('--n_train', default=1e6, type=float)
    p.add_argument('--n_valid', '--n_test', dest='n_valid', default=1e6, type=float)
    p.add_argument('--n_OoD', '--n_W', dest='n_OoD', default=10e6, type=float)
    p.add_argument('--n_sig', '--n_top', dest='n_sig', default=1e6, type=float)
    p.add_argument('--n_const', '--n_constituents', dest='n_const', default=20, type=int)
    p.add_argument('--n_dims', default=3, type=int)
    p.add_argument('--batch_size', default=1e4, type=float)
    p.add_argument('--n_epochs', default=100, type=int)
    p.add_argument('--FC_layers', default=[80, 40, 20, 10], type=int, nargs='+')
    p.add_argument('--FCN_neurons', default=None, type=int, nargs='+')           # README spelling: hidden widths only
    p.add_argument('--latent_dim', default=None, type=int)                        # README spelling: FC_layers[-1]
    p.add_argument('--lr', default=1e-3, type=float)
    p.add_argument('--beta', default=0, type=float)
    p.add_argument('--lamb', default=0, type=float)
    p.add_argument('--margin', default=1, type=float)
    p.add_argument('--n_iter', default=1, type=int)
    p.add_argument('--OE_type', default='KLD')
    p.add_argument('--weight_type', default='X-S')
    p.add_argument('--model_in', default='')
    p.add_argument('--model_out', default='model.h5')
    p.add_argument('--const_scaler_type', default='')
    p.add_argument('--const_scaler_in', default='')
    p.add_argument('--const_scaler_out', default='')
    p.add_argument('--hist_file', default='history.pkl')
    p.add_argument('--output_dir', default='outputs')
    p.add_argument('--plotting', default='OFF')                                   # figures are out of scope here
    p.add_argument('--apply_cuts', '--apply_cut', dest='apply_cuts', default='OFF')
    p.add_argument('--normal_losses', default='ON')
    p.add_argument('--decorrelation', default='OFF')
    p.add_argument('--slurm_id', default=0, type=int)
    p.add_argument('--constituents', default='ON')
    p.add_argument('--HLVs', default='OFF')
    p.add_argument('--patience', default=3, type=int)                             # hard-coded 3 in models.py:222
    p.add_argument('--n_gpus', default=1, type=int)
    p.add_argument('--precision', default='bf16x3', choices=['fp32', 'bf16x3', 'bf16'])
    p.add_argument('--seed', default=1234, type=int)
    a = p.parse_args(argv)
    for key in ['n_train', 'n_valid', 'n_OoD', 'n_sig', 'batch_size']:           # vae.py:50
        vars(a)[key] = int(vars(a)[key])
    if a.FCN_neurons is not None:                                                 # last element = latent dim
        a.FC_layers = list(a.FCN_neurons) + [a.latent_dim if a.latent_dim is not None else 10]
    elif a.latent_dim is not None:
        a.FC_layers = list(a.FC_layers[:-1]) + [a.latent_dim]
    if a.HLVs == 'ON':
        raise SystemExit('--HLVs ON needs the 12 high-level variables of the HDF5 files; synthetic samples carry constituents only')
    if a.const_scaler_out == '':
        a.const_scaler_out = 'const_' + a.const_scaler_type + '.pkl'
    for k in ['model_in', 'model_out', 'const_scaler_in', 'const_scaler_out', 'hist_file']:   # vae.py:58-65
        vars(a)[k] = a.output_dir + '/' + vars(a)[k]
    return a


def _sample(kind, n, args, scaler, seed, device, files=False):
    """load_data (+ apply_scaler, vae.py:125-131,153-162) on synthetic jets -> dict of host arrays.  files=True: the constituents stay
    as the reference's files hold them -- sorted, float16, unscaled (tools/merging.py:31) -- and train_model scales them on the device."""
    import torch
    from . import synthetic as S, utils as U
    raw, jzw = S.make_jets(kind, n, max(args.n_const, 100), seed=seed, device=device)
    if files:
        x, pt, m = U.preprocess_device(raw, args.n_const, args.n_dims, None, True)
        x = x.to(torch.float16)
    else:
        x, pt, m = U.preprocess_device(raw.to(torch.float16), args.n_const, args.n_dims, scaler, True)   # same float16 rounding as the files
    return {'constituents': x.cpu().numpy(), 'weights': np.ones(n, np.float32), 'JZW': jzw.cpu().numpy(),
            'pt': pt.cpu().numpy(), 'm': m.cpu().numpy()}


def main(argv=None):
    args = parse_args(argv)
    import torch
    from . import dist as D, synthetic as S, utils as U, cuts as Cu
    from .models import VariationalAutoEncoder, train_model
    rank, world = D.init()
    if world != args.n_gpus and rank == 0:
        print('note: --n_gpus %d but WORLD_SIZE=%d; launch with torchrun --nproc-per-node N' % (args.n_gpus, world))
    device = torch.device('cuda', int(os.environ.get('LOCAL_RANK', '0')))
    torch.cuda.set_device(device)
    Path(args.output_dir + '/plots').mkdir(parents=True, exist_ok=True)
    input_dim = args.n_dims * args.n_const                                        # vae.py:72-76 (constituents only)
    model = VariationalAutoEncoder(args.FC_layers, input_dim, seed=None if args.n_iter > 1 else 0,
                                   precision=args.precision, max_batch=args.batch_size)
    if args.model_in != args.output_dir + '/' and os.path.isfile(args.model_in):
        model.load_weights(args.model_in)                                          # vae.py:93-102 (no dummy build call needed)
    # scaler (vae.py:103-124): fitted on up to 100k background jets unless a pickled one is given; MaxAbs when no type is asked for
    # (synthetic momenta are O(100 GeV): unscaled inputs would overflow exp(log_var))
    scaler = None
    if os.path.isfile(args.const_scaler_in):
        scaler = pickle.load(open(args.const_scaler_in, 'rb'))
    else:
        raw, _ = S.make_jets('QCD', min(args.n_train, 100_000), max(args.n_const, 100), seed=args.seed, device=device)
        x0, _, _ = U.preprocess_device(raw.to(torch.float16), args.n_const, args.n_dims, None, True, want_4v=False)
        stype = args.const_scaler_type or 'MaxAbsScaler'
        scaler = U.fit_scaler(x0.cpu().numpy(), args.n_dims, args.const_scaler_out if rank == 0 else None, stype)   # utils.py:389-405 (pickles it)
    affine = U._scaler_affine(scaler, input_dim) is not None
    if args.n_epochs > 0:
        # affine scalers: the host keeps the file-format float16 rows and train_model applies the scaler on the device per batch;
        # Quantile / Power transformers: utils.apply_scaler (device interpolation kernels) on the whole sample first, as vae.py:125-131
        def train_sample(kind, n, seed):
            if affine: return _sample(kind, n, args, None, seed, device, files=True)
            smp = _sample(kind, n, args, None, seed, device, files=True)
            smp['constituents'] = U.apply_scaler(smp['constituents'].astype(np.float32), args.n_dims, scaler)
            return smp
        bkg = train_sample('QCD', args.n_train, args.seed + 1)
        ood = train_sample('W', args.n_train, args.seed + 2) if args.lamb != 0 else bkg
        vbkg = train_sample('QCD', min(args.n_valid, args.n_train), args.seed + 3)
        vood = train_sample('W', len(vbkg['weights']), args.seed + 4) if args.lamb != 0 else vbkg
        train_model(model, [(bkg, ood)], [(vbkg, vood)], args.OE_type, args.n_epochs, args.batch_size, args.beta, args.lamb,
                    args.margin, args.lr, args.hist_file, args.model_in, args.model_out, patience=args.patience,
                    scaler=scaler if affine else None)
        D.barrier()
        if os.path.isfile(args.model_out):
            model.load_weights(args.model_out)                                     # vae.py:141: score with the BEST checkpoint, not the last epoch
    # ---- anomaly-scoring pass (vae.py:145-176): QCD + top, sharded across ranks
    n_b, n_s = args.n_valid, min(args.n_sig, args.n_valid)
    lo_b, hi_b = D.shard_bounds(n_b, rank, world); lo_s, hi_s = D.shard_bounds(n_s, rank, world)
    def test_sample(kind, n, seed):
        smp = _sample(kind, n, args, scaler if affine else None, seed, device)
        if not affine: smp['constituents'] = U.apply_scaler(smp['constituents'], args.n_dims, scaler)
        return smp
    parts = [test_sample('QCD', hi_b - lo_b, args.seed + 10 + rank), test_sample('top', hi_s - lo_s, args.seed + 100 + rank)]
    sample = {k: np.concatenate([p_[k] for p_ in parts]) for k in parts[0]}
    y_true = np.where(sample['JZW'] == -1, 0, 1)                                   # vae.py:150
    losses, finite = Cu.score_sample(model, sample['constituents'], args.n_iter, normal_losses='ON' if args.decorrelation == 'ON' else args.normal_losses)
    sample = {k: v[finite] for k, v in sample.items()}; y_true = y_true[finite]     # utils.filtering
    losses = {k: v[finite] for k, v in losses.items()}
    if args.decorrelation == 'ON':                                                  # plots.py:36-39
        losses['MAE'] = Cu.mass_deco(y_true, sample, losses['MAE'], deco='2d')
    if rank == 0:
        for k, v in losses.items():
            print('%-7s mean score: QCD %.4f  top %.4f' % (k, v[y_true == 1].mean(), v[y_true == 0].mean()))
    if args.apply_cuts == 'ON':
        # the test set is sharded: scores are gathered for the threshold search, every rank cuts its own jets, counts are reduced
        res = Cu.generate_cuts(y_true, sample, losses['MAE'], 'MAE')
        if rank == 0:
            pickle.dump(res, open(args.output_dir + '/plots/cut_histograms.pkl', 'wb'))
            for name, r in res.items():
                print('%-16s loss_cut %.5f  n_pass %d' % (name, r['loss_cut'], r['n_pass']))
    D.barrier()
    return 0


if __name__ == '__main__':
    sys.exit(main())
