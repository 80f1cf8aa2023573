"""CPU-side checks of the boundary: the shared library loads, exports every symbol include/oevae.h (the drop-in boundary) and
include/oevae_test.h (kernel-level test / tuning entry points) declare, and the ctypes tables mirror the headers.  No compute calls (there is no GPU on the CPU test box)."""
import os
import re
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header='oevae.h'):
    src = open(os.path.join(ROOT, 'include', header)).read()
    return sorted(set(re.findall(r'OEVAE_API[^;]*?\b(oevae_\w+)\s*\(', src)))


def test_library_builds_and_exports_header_symbols():
    lib = os.path.join(ROOT, 'atlas_vae_b200', 'liboevae.so')
    if not os.path.isfile(lib):
        subprocess.run(['bash', os.path.join(ROOT, 'build.sh')], check=True)
    out = subprocess.run(['nm', '-D', '--defined-only', lib], check=True, capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r'\b(oevae_\w+)\b', out)))
    declared = _declared()
    assert len(declared) >= 25
    assert not any(d.startswith('oevae_probe') or d == 'oevae_gemm' for d in declared)   # test hooks stay out of the public header
    assert exported == sorted(declared + _declared('oevae_test.h'))


def test_ctypes_table_matches_header():
    from atlas_vae_b200 import _ffi
    assert sorted(_ffi.SIGNATURES) == _declared()
    assert sorted(_ffi.TEST_SIGNATURES) == _declared('oevae_test.h')
    L = _ffi.lib()                              # dlopen + resolve every symbol
    assert L.oevae_version() == 200
    assert _ffi.launch_count() == 0


def test_argument_counts_match_header():
    """Each ctypes argtypes list has as many entries as the C prototype has parameters."""
    from atlas_vae_b200 import _ffi
    src = open(os.path.join(ROOT, 'include', 'oevae.h')).read() + open(os.path.join(ROOT, 'include', 'oevae_test.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    for name, (_, argtypes) in list(_ffi.SIGNATURES.items()) + list(_ffi.TEST_SIGNATURES.items()):
        m = re.search(r'\b%s\s*\(([^)]*)\)' % name, src)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ('', 'void') else params.count(',') + 1
        assert n == len(argtypes), (name, n, len(argtypes))


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from atlas_vae_b200._ffi import OevaeError
    from atlas_vae_b200 import VariationalAutoEncoder
    with pytest.raises(OevaeError):
        VariationalAutoEncoder([8, 4], 6, seed=0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'atlas_vae_b200')
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', txt, flags=re.M), f
