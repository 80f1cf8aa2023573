"""mass_deco on the device (OE-VAE/plots.py:54-85) against outputs of the reference's own get_bins / cum_distribution /
mass_deco bodies (tests/golden/make_golden_deco.py -> ref_deco.npz)."""
import os
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('deco', ['2d', 'm', 'pt'])
def test_mass_deco_matches_reference(golden_dir, deco):
    from atlas_vae_b200 import cuts as Cc
    from conftest import record_parity
    g = np.load(os.path.join(golden_dir, 'ref_deco.npz'))
    sample = {'m': g['m'], 'pt': g['pt']}
    got = Cc.mass_deco(g['y_true'], sample, g['loss'].copy(), deco=deco)
    ref = g['deco_' + deco]
    assert got.dtype == np.float32 and got.shape == ref.shape
    same = (got == ref) | (np.isnan(got) & np.isnan(ref))
    record_parity('mass_deco(%s): fraction of jets not bit-identical to the reference' % deco, 1 - same.mean(), 0.0,
                  {'max_abs': float(np.nanmax(np.abs(got.astype(np.float64) - ref)))})
    np.testing.assert_array_equal(got, ref)
    # the decorrelated background score is flat in every cell: its mean is (n + 1) / (2 n) for a cell of n jets, 0.5 for large cells
    b = got[g['y_true'] == 1]
    assert 0.5 <= b.mean() < 0.6


def test_get_bins_matches_reference(golden_dir):
    from atlas_vae_b200 import cuts as Cc
    g = np.load(os.path.join(golden_dir, 'ref_deco.npz'))
    m_t = torch.as_tensor(g['m']).cuda(); keep = torch.as_tensor((g['y_true'] == 1).astype(np.uint8)).cuda()
    np.testing.assert_array_equal(Cc.get_bins_device(m_t, keep, True), g['m_bins'])
