"""ctypes binding of libsbayes_b200.so (C ABI: include/sbayes_b200.h).

There is NO CPU fallback: importing this module without the built library raises, and every compute call
needs a CUDA device (`sb_create` returns SB_E_NOGPU otherwise).
"""
import ctypes as C
import os

import numpy as np

LIB_PATH = os.environ.get('SBAYES_B200_LIB') or os.path.join(os.path.dirname(os.path.abspath(__file__)), 'csrc', 'libsbayes_b200.so')      # override: instrumented builds

SB_N_OPS = 12
SB_F_INHERITANCE, SB_F_SAMPLE_SOURCE, SB_F_WARMUP = 1, 2, 4
SIZE_TYPES = {'none': 0, 'uniform': 1, 'quadratic': 2}
GEO_TYPES = {'uniform': 0, 'cost_based': 1}
DIR_TYPES = {'uniform': 0, 'counts': 1}
OP_IDS = {
    'shrink_zone': 0, 'grow_zone': 1, 'swap_zone': 2,
    'gibbs_sample_sources': 3, 'gibbs_sample_weights': 4, 'gibbs_sample_p_global': 5,
    'gibbs_sample_p_zones': 6, 'gibbs_sample_p_families': 7,
    'alter_weights': 8, 'alter_p_global': 9, 'alter_p_zones': 10, 'alter_p_families': 11,
}
OP_NAMES = {v: k for k, v in OP_IDS.items()}
ERRORS = {-1: 'SB_E_ARG', -2: 'SB_E_CUDA', -3: 'SB_E_STATE', -4: 'SB_E_OVERLAP', -5: 'SB_E_NOGPU', -6: 'SB_E_NUMERIC'}
SB_ERR_NONFINITE, SB_ERR_GEO_CAP, SB_ERR_NAPP = 1, 2, 4

# every symbol include/sbayes_b200.h declares (tests/test_abi_symbols.py checks the .so exports them all)
SYMBOLS = [
    'sb_version', 'sb_device_count', 'sb_create', 'sb_destroy', 'sb_last_error', 'sb_set_stream', 'sb_sync',
    'sb_upload_data', 'sb_upload_locations', 'sb_set_priors', 'sb_set_state', 'sb_set_states', 'sb_get_state', 'sb_copy_state',
    'sb_eval', 'sb_eval_likelihood_async', 'sb_read_ll', 'sb_eval_likelihood_noalloc', 'sb_eval_components', 'sb_source_posterior',
    'sb_counts', 'sb_recount_async', 'sb_device_scalars', 'sb_move_log_q', 'sb_pair_log_q', 'sb_set_operators', 'sb_run', 'sb_set_sample_from_prior',
    'sb_run_events', 'sb_read_event_scalars', 'sb_read_snapshot', 'sb_snapshot_ring', 'sb_read_chain_scalars', 'sb_chain_errors',
    'sb_read_op_stats', 'sb_step_counter', 'sb_kernel_launches', 'sb_rng_sample',
]


class SbDims(C.Structure):
    _fields_ = [('n_sites', C.c_int32), ('n_features', C.c_int32), ('n_states', C.c_int32),
                ('n_zones', C.c_int32), ('n_families', C.c_int32), ('n_chains', C.c_int32),
                ('chain_offset', C.c_int32), ('flags', C.c_uint32)]


class SbPriorCfg(C.Structure):
    _fields_ = [('size_type', C.c_int32), ('geo_type', C.c_int32), ('geo_scale', C.c_double),
                ('p_global_type', C.c_int32), ('p_families_type', C.c_int32)]


class SbOperatorCfg(C.Structure):
    _fields_ = [('n_ops', C.c_int32), ('op_ids', C.c_int32 * SB_N_OPS), ('op_prob', C.c_double * SB_N_OPS),
                ('precision_weights', C.c_double), ('precision_p_global', C.c_double),
                ('precision_p_zones', C.c_double), ('precision_p_families', C.c_double),
                ('min_size', C.c_int32), ('gibbs_feature_fraction', C.c_double), ('seed', C.c_uint64)]


class SbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f'{ERRORS.get(code, code)}: {msg}')
        self.code = code


_lib = None


def load():
    """Load the shared library (fails loudly if it has not been built: `python -c 'import __graft_entry__ as g; g.build()'`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f'{LIB_PATH} is missing: the CUDA library has not been built '
                           f'(run __graft_entry__.build()); there is no CPU fallback')
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_double
    sig = {
        'sb_version': ([], i32), 'sb_device_count': ([], i32),
        'sb_create': ([C.POINTER(vp), i32, C.POINTER(SbDims)], i32),
        'sb_destroy': ([vp], i32), 'sb_last_error': ([vp], C.c_char_p),
        'sb_set_stream': ([vp, vp], i32), 'sb_sync': ([vp], i32),
        'sb_upload_data': ([vp] + [vp] * 6, i32),
        'sb_upload_locations': ([vp, vp], i32),
        'sb_set_priors': ([vp, C.POINTER(SbPriorCfg)] + [vp] * 4, i32),
        'sb_set_state': ([vp, i32] + [vp] * 6, i32),
        'sb_set_states': ([vp, i32, i32] + [vp] * 6, i32),
        'sb_get_state': ([vp, i32] + [vp] * 6, i32),
        'sb_copy_state': ([vp, i32, i32], i32),
        'sb_eval': ([vp, i32, i32, i32, vp, vp, vp], i32),
        'sb_eval_likelihood_async': ([vp, i32, i32, i32], i32),
        'sb_read_ll': ([vp, i32, i32, vp], i32),
        'sb_eval_likelihood_noalloc': ([vp, i32, i32, vp], i32),
        'sb_eval_components': ([vp, i32, vp, vp], i32),
        'sb_source_posterior': ([vp, i32, vp], i32),
        'sb_counts': ([vp, i32, i32] + [vp] * 5, i32),
        'sb_move_log_q': ([vp, i32, i32, i32, i32, i32, dbl, C.POINTER(dbl), C.POINTER(dbl)], i32),
        'sb_set_operators': ([vp, C.POINTER(SbOperatorCfg), vp, vp], i32),
        'sb_run': ([vp, i64], i32),
        'sb_pair_log_q': ([vp, dbl, dbl, dbl, dbl, dbl, C.POINTER(dbl)], i32),
        'sb_set_sample_from_prior': ([vp, i32], i32),
        'sb_run_events': ([vp, i64, i32, vp, vp, i32], i32),
        'sb_read_event_scalars': ([vp, i32, vp, vp], i32),
        'sb_read_snapshot': ([vp, i32, i32] + [vp] * 5, i32),
        'sb_snapshot_ring': ([vp, C.POINTER(vp), C.POINTER(i64), C.POINTER(i64), vp, vp], i32),
        'sb_chain_errors': ([vp, vp], i32),
        'sb_recount_async': ([vp, i32, i32], i32),
        'sb_device_scalars': ([vp, C.POINTER(vp), C.POINTER(vp)], i32),
        'sb_read_chain_scalars': ([vp, vp, vp], i32),
        'sb_read_op_stats': ([vp, vp, vp], i32),
        'sb_step_counter': ([vp, C.POINTER(i64)], i32),
        'sb_kernel_launches': ([vp, C.POINTER(i64)], i32),
        'sb_rng_sample': ([vp, i32, dbl, C.c_uint64, i32, vp], i32),
    }
    for name, (argtypes, restype) in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype
    _lib = lib
    return lib


def ptr(a):
    """Host pointer of a C-contiguous numpy array (or NULL)."""
    if a is None:
        return None
    assert a.flags['C_CONTIGUOUS']
    return a.ctypes.data_as(C.c_void_p)


def as_c(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)
