"""ctypes binding of liboevae.so (include/oevae.h).

The product path has NO fallback: if the shared library is missing or a call fails, an exception is raised.
torch is used only as the device allocator / stream provider (tensor.data_ptr(), current stream).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'liboevae.so')

OE_TYPES = {'KLD': 0, 'MSE': 1, 'MAE': 2, 'MSE-margin': 3, 'MAE-margin': 4}
PRECISIONS = {'fp32': 0, 'bf16x3': 1, 'bf16': 3}
NOISE_MODES = {'fixed': 0, 'per_step': 1, 'external': 2}


class OevaeError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [('input_dim', C.c_int32), ('n_hidden', C.c_int32), ('hidden', C.c_int32 * 8),
                ('latent_dim', C.c_int32), ('oe_type', C.c_int32), ('precision', C.c_int32),
                ('noise_mode', C.c_int32), ('device', C.c_int32), ('seed', C.c_uint64), ('max_batch', C.c_int64)]


class HostLoad(C.Structure):
    """oevae_host_load (include/oevae.h)"""
    _fields_ = [('x_host', C.c_void_p), ('w_host', C.c_void_p), ('xo_host', C.c_void_p), ('wo_host', C.c_void_p),
                ('x_dtype', C.c_int32), ('update', C.c_int32),
                ('n_rows', C.c_int64), ('batch_size', C.c_int64), ('first_batch', C.c_int64), ('n_batches', C.c_int64),
                ('rank', C.c_int32), ('world', C.c_int32), ('noise_row0', C.c_int64),
                ('center', C.c_void_p), ('scale', C.c_void_p),
                ('x_resident', C.c_void_p), ('w_resident', C.c_void_p), ('xo_resident', C.c_void_p), ('wo_resident', C.c_void_p),
                ('resident_stride', C.c_int64),
                ('lr', C.c_float), ('beta', C.c_float), ('lamb', C.c_float), ('margin', C.c_float),
                ('sums_acc', C.c_void_p), ('sums_trace_host', C.c_void_p),
                ('h2d_bytes', C.c_int64)]


_lib = None
_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

# name -> (restype, argtypes); mirrors include/oevae.h one to one (TEST_SIGNATURES: include/oevae_test.h)
SIGNATURES = {
    'oevae_create': (C.c_int, [C.POINTER(Config), C.POINTER(_vp)]),
    'oevae_destroy': (C.c_int, [_vp]),
    'oevae_version': (C.c_int, []),
    'oevae_last_error': (C.c_char_p, [_vp]),
    'oevae_n_params': (_i64, [_vp]),
    'oevae_param_layout': (C.c_int, [_vp, C.POINTER(_i64), C.POINTER(_i32), C.POINTER(_i32), _i32]),
    'oevae_set_params_host': (C.c_int, [_vp, _vp, _i64]),
    'oevae_get_params_host': (C.c_int, [_vp, _vp, _i64]),
    'oevae_set_opt_state_host': (C.c_int, [_vp, _vp, _vp, _i64, _i64]),
    'oevae_get_opt_state_host': (C.c_int, [_vp, _vp, _vp, _i64, C.POINTER(_i64)]),
    'oevae_params_device': (_vp, [_vp]),
    'oevae_reset_opt_state': (C.c_int, [_vp]),
    'oevae_set_noise_offset': (C.c_int, [_vp, C.c_uint64]),
    'oevae_set_oe_type': (C.c_int, [_vp, _i32]),
    'oevae_train_step': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _f32, _f32, _f32, _f32, _vp, _vp, _vp]),
    'oevae_loss_and_grad': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _f32, _f32, _f32, _vp, _vp, _vp, _vp]),
    'oevae_apply_update': (C.c_int, [_vp, _vp, _f32, _vp]),
    'oevae_train_batches_host': (C.c_int, [_vp, C.POINTER(HostLoad), _vp]),
    'oevae_dp_export': (C.c_int, [_vp, _vp]),
    'oevae_dp_connect': (C.c_int, [_vp, _i32, _i32, _vp]),
    'oevae_dp_last_grad_host': (C.c_int, [_vp, _vp, _i64]),
    'oevae_train_step_dp': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _f32, _f32, _f32, _f32, _vp, _vp, _vp]),
    'oevae_eval_losses': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _f32, _f32, _f32, _vp, _vp, _vp]),
    'oevae_encode': (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    'oevae_score': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'oevae_preprocess': (C.c_int, [_vp, _i32, _i64, _i32, _i32, _i32, _vp, _vp, _i32, _vp, _vp, _vp, _i32, _vp]),
    'oevae_affine': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _vp]),
    'oevae_ingest_f16': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _i32, _vp]),
    'oevae_quantile_normal': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _i32, C.c_double, C.c_double, _i32, _vp]),
    'oevae_power_yj': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp, _i32, _vp]),
    'oevae_pair_metrics': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp]),
    'oevae_minmax': (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp]),
    'oevae_loss_map': (C.c_int, [_vp, _i64, _i32, _i32, _vp]),
    'oevae_hist': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _f32, _vp, _i32, _f32, _i32, _vp, _vp, _i32, _vp]),
    'oevae_wsum': (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp]),
    'oevae_digitize_count': (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _i32, _vp]),
    'oevae_mass_deco': (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _i32, _vp, _vp, _i32, _vp]),
    'oevae_argsort_f32': (C.c_int, [_vp, _i64, _vp, _i32, _vp]),
    'oevae_argsort_f32_static': (C.c_int, [_vp, _i64, _vp, _i32, _vp]),
    'oevae_ood_pairing': (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, C.c_uint64, _i32, _vp, _vp, _i32, _vp]),
    'oevae_roc': (C.c_int, [_vp, _vp, _i32, _vp, _i64, _vp, _vp, _vp, C.POINTER(_i64), _i32, _vp]),
    'oevae_roc_cuts': (C.c_int, [_vp, _vp, _vp, _i64, _vp, _i32, C.c_double, _vp, _i32, _vp]),
    'oevae_roc_searchsorted': (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _vp, _i32, _vp]),
    'oevae_host_register': (C.c_int, [_vp, _i64]),
    'oevae_host_unregister': (C.c_int, [_vp]),
    'oevae_launch_count': (_i64, []),
    'oevae_profile_enable': (C.c_int, [_i32]),
    'oevae_profile_report': (C.c_int, [C.c_char_p, _i32]),
}
TEST_SIGNATURES = {
    'oevae_dp_stamps_host': (C.c_int, [_vp, _vp, _i64]),
    'oevae_gemm': (C.c_int, [_vp, _i32, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _i32, _i32, _i32, _i32, _vp, _i32, _vp]),
    'oevae_probe_mn': (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _i32, _vp]),
    'oevae_probe_latency': (C.c_int, [_i32, _i32, C.POINTER(_i64), _i32]),
    'oevae_probe_issue': (C.c_int, [_i32, _i32, _i32, _i32, _i32, C.POINTER(_i64), _i32]),
    'oevae_probe_bulk': (C.c_int, [_i32, _i32, _i32, _i32, C.POINTER(_i64), _i32]),
}


def lib():
    """Load liboevae.so (built in-tree by build.sh / __graft_entry__.build()); raises if it is absent."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise OevaeError('%s not found: run ./build.sh (there is no CPU fallback for the OE-VAE hot path)' % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in list(SIGNATURES.items()) + list(TEST_SIGNATURES.items()):
            fn = getattr(L, name)            # AttributeError if the library does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc, handle=None):
    if rc != 0:
        msg = lib().oevae_last_error(handle)
        raise OevaeError('liboevae error %d: %s' % (rc, msg.decode() if msg else '?'))


def ptr(t):
    """device (or host) address of a torch tensor / numpy array, None -> NULL."""
    if t is None:
        return None
    if hasattr(t, 'data_ptr'):
        assert t.is_contiguous(), 'non-contiguous tensor passed to liboevae'
        return t.data_ptr()
    assert t.flags['C_CONTIGUOUS']
    return t.ctypes.data


def launch_count():
    return int(lib().oevae_launch_count())


def profile(on):
    lib().oevae_profile_enable(1 if on else 0)


def profile_report():
    """[(kernel name, launches, total ms)] of the launches made while profiling was on."""
    buf = C.create_string_buffer(1 << 16)
    n = lib().oevae_profile_report(buf, len(buf))
    rows = []
    for line in buf.value.decode().splitlines():
        name, k, ms = line.rsplit(',', 2)
        rows.append((name, int(k), float(ms)))
    return rows
