"""ctypes binding of ``libremixt_b200.so`` (C-ABI declared in ``include/remixt_b200.h``).

There is no fallback: if the shared library is missing the import fails, and if no CUDA
device is present the first ``rmx_create`` fails with a RuntimeError.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# REMIXT_B200_LIB names another build of the same library (A/B measurements of a kernel change on one GPU box)
LIB_PATH = os.environ.get("REMIXT_B200_LIB") or os.path.join(_HERE, "libremixt_b200.so")

RMX_OK, RMX_ERR_CUDA, RMX_ERR_VALUE, RMX_ERR_ASSERT, RMX_ERR_UNSUPPORTED, RMX_ERR_ARG = range(6)

# rmx_array_id
A_H, A_P_BREAKPOINT, A_P_ALLELE_SWAP, A_P_OUTLIER_TOTAL, A_P_OUTLIER_ALLELE = 0, 1, 2, 3, 4
A_POSTERIOR_MARGINALS, A_FRAMELOGPROB, A_TOTAL_LIKELIHOOD_MASK, A_ALLELE_LIKELIHOOD_MASK = 5, 6, 7, 8
A_LOG_TRANSMAT, A_CACHED_LOG_TRANSMAT, A_JOINT_POSTERIOR_MARGINALS = 9, 10, 11
A_L, A_X, A_Y, A_STATE_PATH = 12, 13, 14, 15
A_LL_TABLE = 16
DESC_SHARD = 1

# rmx_scalar_id
SCALAR_IDS = {
    'negbin_r_0': 0, 'negbin_r_1': 1, 'negbin_hdel_mu': 2, 'negbin_hdel_r_0': 3, 'negbin_hdel_r_1': 4,
    'betabin_M_0': 5, 'betabin_M_1': 6, 'betabin_loh_p': 7, 'betabin_loh_M_0': 8, 'betabin_loh_M_1': 9,
    'prior_outlier_total': 10, 'prior_outlier_allele': 11, 'transition_penalty': 12,
    'divergence_weight': 13, 'transition_model': 14, 'hmm_log_norm_const': 15,
}

KERNEL_CLASSES = ['chain', 'lltable', 'frame', 'qsums', 'post', 'hist', 'brk_update', 'elbo_rows', 'pairsum', 'ell', 'round']

# rmx_update_id
U_ALLELE_SWAP, U_CN, U_BREAKPOINT, U_OUTLIER_TOTAL, U_OUTLIER_ALLELE, U_FRAMELOGPROB = range(6)

# every symbol include/remixt_b200.h declares
EXPORTS = [
    'rmx_create', 'rmx_batch_create', 'rmx_batch_view', 'rmx_num_fits', 'rmx_destroy', 'rmx_release_cached_memory', 'rmx_last_error',
    'rmx_device_count',
    'rmx_set_array', 'rmx_get_array', 'rmx_set_scalar', 'rmx_get_scalar', 'rmx_get_derived',
    'rmx_update', 'rmx_variational_update', 'rmx_breakpoint_logq', 'rmx_breakpoint_apply_logq',
    'rmx_comm_load', 'rmx_comm_unique_id', 'rmx_comm_init', 'rmx_comm_destroy',
    'rmx_variational_energy', 'rmx_variational_entropy', 'rmx_elbo',
    'rmx_expected_log_likelihood', 'rmx_expected_log_likelihood_partial_h',
    'rmx_expected_log_likelihood_grid',
    'rmx_mstep_stage_sample', 'rmx_mstep_stage_rows', 'rmx_mstep_evaluate', 'rmx_mstep_param_search', 'rmx_set_scalar_array', 'rmx_host_sample_segments',
    'rmx_infer_cn', 'rmx_calculate_log_transmat', 'rmx_eval',
    'rmx_sum_product', 'rmx_sum_product_2paramtrans', 'rmx_max_product', 'rmx_max_product_noalloc',
    'rmx_launch_count', 'rmx_global_launch_count', 'rmx_kernel_timing', 'rmx_chain_timing', 'rmx_synchronize', 'rmx_stream',
    'rmx_event_record', 'rmx_event_elapsed', 'rmx_fp64_peak',
]


class RmxDesc(C.Structure):
    _fields_ = [
        ('num_clones', C.c_int32), ('num_segments', C.c_int32), ('num_breakpoints', C.c_int32),
        ('normal_contamination', C.c_int32), ('num_cn_states', C.c_int32), ('num_brk_states', C.c_int32),
        ('device', C.c_int32), ('flags', C.c_int32),
        ('cn_states', C.c_void_p), ('cn_table', C.c_void_p), ('normal_copies', C.c_void_p),
        ('brk_states', C.c_void_p), ('h_init', C.c_void_p), ('l', C.c_void_p), ('x', C.c_void_p),
        ('y', C.c_void_p), ('is_telomere', C.c_void_p), ('breakpoint_idx', C.c_void_p),
        ('breakpoint_orient', C.c_void_p),
        ('transition_penalty', C.c_double), ('divergence_weight', C.c_double),
    ]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "remixt_b200: %s is missing; build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C remixt_b200/csrc` (there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    H, I, I64, D, P = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_void_p
    sigs = {
        'rmx_create': [C.POINTER(RmxDesc), C.POINTER(H)],
        'rmx_batch_create': [C.POINTER(RmxDesc), I, P, P, C.POINTER(H)],
        'rmx_batch_view': [H, I, I, C.POINTER(H)],
        'rmx_num_fits': [H, C.POINTER(I)],
        'rmx_destroy': [H],
        'rmx_breakpoint_logq': [H, P],
        'rmx_breakpoint_apply_logq': [H, P],
        'rmx_release_cached_memory': [I],
        'rmx_comm_load': [C.c_char_p],
        'rmx_comm_unique_id': [P],
        'rmx_comm_init': [H, P, I, I],
        'rmx_comm_destroy': [H],
        'rmx_device_count': [C.POINTER(I)],
        'rmx_set_array': [H, I, P, I64],
        'rmx_get_array': [H, I, P, I64],
        'rmx_set_scalar': [H, I, D],
        'rmx_get_scalar': [H, I, P],
        'rmx_get_derived': [H, I, P, I64],
        'rmx_update': [H, I],
        'rmx_variational_update': [H, I],
        'rmx_variational_energy': [H, P],
        'rmx_variational_entropy': [H, P],
        'rmx_elbo': [H, P],
        'rmx_expected_log_likelihood': [H, P, P],
        'rmx_expected_log_likelihood_partial_h': [H, P, P],
        'rmx_expected_log_likelihood_grid': [H, P, I, P, I, P],
        'rmx_mstep_stage_sample': [H, P],
        'rmx_mstep_stage_rows': [H, P, P, I],
        'rmx_mstep_evaluate': [H, P, P, I, P, I, P, P],
        'rmx_mstep_param_search': [H, P, I, D, D, P, I, P, P, P, P, I],
        'rmx_set_scalar_array': [H, I, P, P],
        'rmx_host_sample_segments': [I, I, I, I, P, P, P, P, P, P, I],
        'rmx_infer_cn': [H, P],
        'rmx_calculate_log_transmat': [H, P],
        'rmx_eval': [H, I, I, I, I, I, I, P],
        'rmx_sum_product': [I, I64, I64, P, P, P, P],
        'rmx_sum_product_2paramtrans': [I, I64, I64, P, P, P, D, D],
        'rmx_max_product': [I, I64, I64, P, P, P, C.POINTER(D)],
        'rmx_max_product_noalloc': [I, P, I64, I64, P, P, P, P],
        'rmx_launch_count': [H, C.POINTER(I64)],
        'rmx_global_launch_count': [C.POINTER(I64)],
        'rmx_kernel_timing': [H, I, P, P],
        'rmx_chain_timing': [H, I, C.POINTER(D), C.POINTER(I64)],
        'rmx_synchronize': [H],
        'rmx_stream': [H, C.POINTER(P)],
        'rmx_event_record': [H, I],
        'rmx_event_elapsed': [H, I, H, I, C.POINTER(D)],
        'rmx_fp64_peak': [I, C.POINTER(D)],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    lib.rmx_last_error.argtypes = [H]
    lib.rmx_last_error.restype = C.c_char_p
    return lib


lib = _load()


def last_error(handle=None):
    msg = lib.rmx_last_error(handle)
    return msg.decode('utf-8', 'replace') if msg else ''


def check(rc, handle=None):
    """Map an rmx_status to the exception class the reference raises (bpmodel.pyx `except *`)."""
    if rc == RMX_OK:
        return
    msg = last_error(handle)
    if rc == RMX_ERR_VALUE or rc == RMX_ERR_UNSUPPORTED:
        raise ValueError(msg)
    if rc == RMX_ERR_ASSERT:
        raise AssertionError(msg)
    if rc == RMX_ERR_ARG:
        raise TypeError(msg)
    raise RuntimeError('remixt_b200: ' + msg)


def device_count():
    n = C.c_int(0)
    lib.rmx_device_count(C.byref(n))
    return n.value


def release_cached_memory(device=-1):
    """Return the device blocks kept from destroyed batches to the driver."""
    lib.rmx_release_cached_memory(int(device))


def global_launch_count():
    n = C.c_int64(0)
    lib.rmx_global_launch_count(C.byref(n))
    return n.value


def fp64_peak(device=0):
    """Measured DFMA peak of the device in TFLOP/s."""
    v = C.c_double(0.)
    check(lib.rmx_fp64_peak(int(device), C.byref(v)), None)
    return v.value
