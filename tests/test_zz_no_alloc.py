"""oevae_argsort_f32_static: the allocation-free variant of oevae_argsort_f32 (include/oevae.h; no counterpart in the reference).
Its radix-sort workspace is a static __device__ array of the library, shared by every call, so
the checks are: same order as the allocating entry point (bit-exact: index work) on two calls in a row whose sizes differ -- the second
one runs on the leftovers of the first --, the stated maximum is accepted, anything above it is refused with a message naming the limit,
and the calls leave torch's allocator untouched.  Collected last on purpose."""
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

STATIC_MAX = 262144                                   # OEVAE_ARGSORT_STATIC_MAX


def _both(v):
    from atlas_vae_b200 import _ffi
    from atlas_vae_b200._ffi import ptr, check
    lib = _ffi.lib()
    n = len(v)
    t = torch.as_tensor(v).cuda()
    ref = torch.empty(n, dtype=torch.int64, device='cuda')
    got = torch.full((n,), -1, dtype=torch.int64, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream
    check(lib.oevae_argsort_f32(ptr(t), n, ptr(ref), 0, stream))
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()['allocation.all.allocated']
    check(lib.oevae_argsort_f32_static(ptr(t), n, ptr(got), 0, stream))
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()['allocation.all.allocated'] == before
    return ref.cpu().numpy(), got.cpu().numpy()


def _values(n, seed, ties):
    rng = np.random.default_rng(seed)
    v = rng.gamma(3.0, 30.0, n).astype(np.float32)
    v[rng.uniform(size=n) < 0.1] *= -1                # both signs
    if ties: v = np.round(v, 0).astype(np.float32)    # long runs of ties: the stable order matters
    return v


def test_static_argsort_equals_allocating_argsort_on_two_calls_in_a_row():
    # first call: a large, tie-heavy input fills the whole workspace; second call: a short, ragged one runs on its leftovers
    for n, seed, ties in ((200_003, 1, True), (4_097, 2, False), (STATIC_MAX, 3, True), (1, 4, False)):
        v = _values(n, seed, ties)
        ref, got = _both(v)
        np.testing.assert_array_equal(got, ref)
        np.testing.assert_array_equal(got, np.argsort(v, kind='stable'))


def test_static_argsort_refuses_sizes_beyond_the_maximum():
    from atlas_vae_b200 import _ffi
    from atlas_vae_b200._ffi import ptr
    lib = _ffi.lib()
    n = STATIC_MAX + 1
    t = torch.zeros(n, dtype=torch.float32, device='cuda')
    out = torch.full((n,), -7, dtype=torch.int64, device='cuda')
    rc = lib.oevae_argsort_f32_static(ptr(t), n, ptr(out), 0, torch.cuda.current_stream().cuda_stream)
    assert rc < 0
    msg = lib.oevae_last_error(None).decode()
    assert 'OEVAE_ARGSORT_STATIC_MAX' in msg and str(STATIC_MAX) in msg, msg
    torch.cuda.synchronize()
    assert bool((out == -7).all())                    # nothing was launched
    # and the entry point still works afterwards
    v = _values(10_000, 5, True)
    ref, got = _both(v)
    np.testing.assert_array_equal(got, ref)
