"""Data parallel over NVLink peer memory (dp_kernels.cuh): needs >= 2 GPUs; skipped on a one-GPU box."""
import os, subprocess, sys
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu


def test_peer_memory_update_matches_allreduce_and_single_rank():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip('needs 2 GPUs')
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29533', os.path.join(here, 'dp_worker.py')]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert 'dp_worker ok' in r.stdout
