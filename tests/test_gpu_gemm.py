"""Kernel-level test of the layered Dense contraction (oevae_gemm): CUDA-core fp32 and the tcgen05 paths against
float64 NumPy for the three operand orientations (forward, dX, dW) and ragged / tile-crossing shapes."""
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

SHAPES = [(5, 7, 3), (128, 64, 16), (300, 400, 80), (1000, 256, 400), (401, 1100, 256), (129, 1, 5), (64, 200, 130), (1, 129, 1)]
BAR = {'fp32': 5e-6, 'bf16x3': 1e-4, 'bf16': 2e-2}          # max |err| / max |ref|


@pytest.mark.parametrize('precision', ['fp32', 'bf16x3', 'bf16'])
def test_gemm_orientations(precision):
    from atlas_vae_b200.engine import DeviceVAE
    rng = np.random.default_rng(0)
    dev = DeviceVAE([8, 4], 6, 'KLD', precision, 'external', 0)
    for (R, T, Cn) in SHAPES:
        A = rng.normal(size=(R, T)); Bm = rng.normal(size=(T, Cn)); bias = rng.normal(size=Cn)
        ref = A @ Bm
        for kind, (a, b) in enumerate([(A, Bm), (A, Bm.T.copy()), (A.T.copy(), Bm)]):
            out = dev.gemm(kind, a.astype(np.float32), b.astype(np.float32)).cpu().numpy()
            err = np.abs(out - ref).max() / np.abs(ref).max()
            assert err < BAR[precision], (precision, kind, (R, T, Cn), err)
        out = dev.gemm(0, A.astype(np.float32), Bm.astype(np.float32), bias.astype(np.float32), relu=True).cpu().numpy()
        ref2 = np.maximum(ref + bias, 0)
        assert np.abs(out - ref2).max() / max(np.abs(ref2).max(), np.abs(ref).max()) < BAR[precision]
