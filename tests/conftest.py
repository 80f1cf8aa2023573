import os, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def golden_dir():
    return GOLDEN


def record_parity(what, achieved_rel, bound, extra=None):
    """Append {test, what, achieved_rel, bound} to gpurun_out/r02_parity.jsonl (the GPU box brings gpurun_out/ back; the file is
    summarised into profiles/r02_parity.json).  Every _close* helper of the GPU tests calls this, so a regression from 1e-6 to 1e-3
    that still passes its bound is visible in the record."""
    import json
    test = os.environ.get('PYTEST_CURRENT_TEST', '?').split(' ')[0]
    row = {'test': test, 'what': what, 'achieved_rel': float(achieved_rel), 'bound': float(bound)}
    if extra: row.update(extra)
    try:
        d = os.path.join(ROOT, 'gpurun_out')
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, 'r02_parity.jsonl'), 'a') as f:
            f.write(json.dumps(row) + '\n')
    except OSError:
        pass
