"""NumPy restatement of the constituent preprocessing feeding the device buffers (TEST INFRASTRUCTURE).

  jets_sorting   utils.py:564-571   -> sort_constituents()
  load_data      utils.py:89-97     -> select_constituents()   (keep first n_const, zero-pad)
  load_data      utils.py:115-119   -> drop_energy()           (n_dims == 3)
  get_4v         utils.py:300-308   -> jet_pt_mass()
  apply_scaling  utils.py:406-417   -> apply_affine()          (MaxAbs / Robust scalers are affine per column)

PINNED against the reference's own jets_sorting / get_4v bodies (tests/golden/ref_preprocess.npz).
"""
import numpy as np


def sort_constituents(jets):
    """pT-descending constituent order; ties resolved exactly as np.argsort(pt)[:, ::-1] (quicksort, reversed)."""
    jets = np.reshape(np.float32(jets), (-1, jets.shape[1] // 4, 4))
    pt = np.sqrt(jets[:, :, 1] ** 2 + jets[:, :, 2] ** 2)
    order = np.argsort(pt, axis=-1)[:, ::-1]
    jets = np.take_along_axis(jets, order[:, :, None], axis=1)
    return jets.reshape(len(jets), -1)


def select_constituents(sorted_jets, n_const):
    out = sorted_jets[:, :4 * n_const]
    if out.shape[1] < 4 * n_const:
        out = np.hstack([out, np.zeros((len(out), 4 * n_const - out.shape[1]), np.float32)])
    return out


def drop_energy(x4):
    n = x4.shape[1] // 4
    return x4.reshape(-1, n, 4)[..., 1:].reshape(len(x4), -1)


def jet_pt_mass(x4):
    s = np.sum(np.float32(x4).reshape(len(x4), -1, 4), axis=1)
    E, px, py, pz = s[:, 0], s[:, 1], s[:, 2], s[:, 3]
    pt = np.sqrt(px ** 2 + py ** 2)
    m = np.sqrt(np.maximum(0, E ** 2 - px ** 2 - py ** 2 - pz ** 2))
    return pt, m


def apply_affine(x, center, scale):
    """(x - center) / scale per column in fp32: what sklearn's MaxAbsScaler (center 0) / RobustScaler transform do."""
    x = np.float32(x)
    if center is not None:
        x = x - np.float32(center)
    return x / np.float32(scale)


def preprocess(raw, n_const, n_dims=4, center=None, scale=None):
    """raw [N, 4*n_raw] (E,px,py,pz per constituent) -> (X [N, n_dims*n_const] fp32, pt [N], m [N])."""
    x4 = select_constituents(sort_constituents(raw), n_const)
    pt, m = jet_pt_mass(x4)
    x = drop_energy(x4) if n_dims == 3 else x4
    if scale is not None:
        x = apply_affine(x, center, scale)
    return np.ascontiguousarray(x, np.float32), pt, m
