"""ctypes binding of libmopo_b200.so (the C ABI in include/mopo_b200.h).

There is NO CPU fallback: if the CUDA library is missing or fails to load, importing this module
raises.  torch is used only for device memory, streams and torch.distributed.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('MOPO_B200_LIB') or os.path.join(_HERE, 'libmopo_b200.so')   # override: developer builds (watchdog)

OK, EINVAL, ESTATE, ECUDA, EUNSUPPORTED = 0, -1, -2, -3, -4
TERM_NEVER, TERM_HOPPER, TERM_WALKER2D, TERM_ANT, TERM_HUMANOID = range(5)
PEN_NONE, PEN_LEARNED_VAR, PEN_DISAGREEMENT = range(3)
PREC_FP32, PREC_BF16, PREC_FP32_REF = 0, 1, 2
F_ACTIONS, F_REWARDS, F_TERMINALS, F_OBS, F_NEXT_OBS = range(5)
IMG_W, IMG_B, IMG_MU, IMG_SIGMA, IMG_MAX_LOGVAR, IMG_MIN_LOGVAR = range(6)

c_f32p = C.c_void_p  # device/host float*, passed as raw addresses


class EnsCfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ('obs_dim', 'act_dim', 'rew_dim', 'hidden_dim', 'n_hidden',
                                          'num_networks', 'num_elites', 'separate_mean_var', 'device')]


class StepArgs(C.Structure):
    _fields_ = [('obs', C.c_void_p), ('act', C.c_void_p), ('B', C.c_int64),
                ('term_kind', C.c_int32), ('penalty_kind', C.c_int32), ('penalty_coeff', C.c_float),
                ('deterministic', C.c_int32), ('precision', C.c_int32),
                ('noise', C.c_void_p), ('model_inds', C.c_void_p), ('seed', C.c_uint64), ('step_id', C.c_uint64),
                ('origin', C.c_void_p),
                ('next_obs', C.c_void_p), ('rewards', C.c_void_p), ('terminals', C.c_void_p),
                ('penalty', C.c_void_p), ('unpenalized', C.c_void_p), ('model_means', C.c_void_p),
                ('model_stds', C.c_void_p), ('log_prob', C.c_void_p), ('dev', C.c_void_p),
                ('model_inds_out', C.c_void_p)]


class RolloutArgs(C.Structure):
    _fields_ = [('start_obs', C.c_void_p), ('B', C.c_int64), ('rollout_length', C.c_int32),
                ('actions', C.POINTER(C.c_void_p)),
                ('term_kind', C.c_int32), ('penalty_kind', C.c_int32), ('penalty_coeff', C.c_float),
                ('deterministic', C.c_int32), ('precision', C.c_int32),
                ('noise', C.POINTER(C.c_void_p)), ('model_inds', C.POINTER(C.c_void_p)),
                ('seed', C.c_uint64), ('rollout_id', C.c_uint64), ('origin_base', C.c_int64),
                ('steps_added', C.POINTER(C.c_int64)), ('action_events', C.POINTER(C.c_void_p)),
                ('policy', C.c_void_p), ('rollout_random', C.c_int32), ('policy_deterministic', C.c_int32)]


class PolicyArgs(C.Structure):
    _fields_ = [('obs', C.c_void_p), ('B', C.c_int64), ('deterministic', C.c_int32), ('random_uniform', C.c_int32),
                ('precision', C.c_int32), ('noise', C.c_void_p), ('seed', C.c_uint64), ('step_id', C.c_uint64),
                ('origin', C.c_void_p), ('origin_base', C.c_int64), ('actions', C.c_void_p), ('mu_out', C.c_void_p),
                ('std_out', C.c_void_p)]


class SacHParams(C.Structure):
    _fields_ = [('reward_scale', C.c_float), ('discount', C.c_float), ('tau', C.c_float), ('policy_lr', C.c_float),
                ('q_lr', C.c_float), ('target_entropy', C.c_float), ('target_update_interval', C.c_int32), ('reserved', C.c_int32)]


class SacBatch(C.Structure):
    _fields_ = [('obs', C.c_void_p), ('act', C.c_void_p), ('rew', C.c_void_p), ('term', C.c_void_p), ('next_obs', C.c_void_p),
                ('n', C.c_int64), ('eps', C.c_void_p), ('seed', C.c_uint64), ('losses_out', C.c_void_p)]


# every symbol include/mopo_b200.h declares: (restype, argtypes)
SYMBOLS = {
    'mopo_last_error': (C.c_char_p, []),
    'mopo_version': (C.c_int, []),
    'mopo_build_hash': (C.c_char_p, []),
    'mopo_launch_count': (C.c_int64, []),
    'mopo_ens_create': (C.c_int, [C.POINTER(EnsCfg), C.POINTER(C.c_void_p)]),
    'mopo_ens_destroy': (C.c_int, [C.c_void_p]),
    'mopo_ens_set_params': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    'mopo_ens_set_elites': (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int32]),
    'mopo_ens_param_image': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    'mopo_ens_refresh_images': (C.c_int, [C.c_void_p, C.c_void_p]),
    'mopo_ens_predict': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    'mopo_ens_predict3d': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    'mopo_ens_predict_noalloc': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mopo_ens_forward': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    'mopo_ens_train_begin': (C.c_int, [C.c_void_p, C.c_float, C.POINTER(C.c_float)]),
    'mopo_ens_train_reset': (C.c_int, [C.c_void_p]),
    'mopo_ens_train_set_deterministic': (C.c_int, [C.c_void_p, C.c_int32]),
    'mopo_ens_train_batches': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                                         C.c_int32, C.c_void_p, C.c_void_p]),
    'mopo_ens_train_mse': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]),
    'mopo_ens_train_snapshot': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p]),
    'mopo_ens_train_restore': (C.c_int, [C.c_void_p, C.c_void_p]),
    'mopo_ens_train_end': (C.c_int, [C.c_void_p, C.c_void_p]),
    'mopo_ens_get_params': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p]),
    'mopo_env_step': (C.c_int, [C.c_void_p, C.POINTER(StepArgs), C.c_void_p]),
    'mopo_termination_fn': (C.c_int, [C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    'mopo_pool_create': (C.c_int, [C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]),
    'mopo_pool_destroy': (C.c_int, [C.c_void_p]),
    'mopo_pool_size': (C.c_int64, [C.c_void_p]),
    'mopo_pool_pointer': (C.c_int64, [C.c_void_p]),
    'mopo_pool_max_size': (C.c_int64, [C.c_void_p]),
    'mopo_pool_field': (C.c_void_p, [C.c_void_p, C.c_int32]),
    'mopo_pool_advance': (C.c_int, [C.c_void_p, C.c_int64]),
    'mopo_pool_add': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    'mopo_pool_gather': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mopo_pool_sample': (C.c_int, [C.c_void_p, C.c_int64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mopo_kernel_timing': (C.c_int, [C.c_void_p, C.c_int]),
    'mopo_kernel_time': (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    'mopo_rollout': (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(RolloutArgs), C.c_void_p]),
    'mopo_rollout_host': (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(RolloutArgs), C.c_void_p]),
    'mopo_sac_create': (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(SacHParams), C.POINTER(C.c_void_p)]),
    'mopo_sac_destroy': (C.c_int, [C.c_void_p]),
    'mopo_sac_set_params': (C.c_int, [C.c_void_p] + [C.POINTER(C.c_void_p)] * 6 + [C.c_float, C.c_void_p]),
    'mopo_sac_get_params': (C.c_int, [C.c_void_p] + [C.POINTER(C.c_void_p)] * 10 + [C.POINTER(C.c_float), C.c_void_p]),
    'mopo_sac_set_targets': (C.c_int, [C.c_void_p] + [C.POINTER(C.c_void_p)] * 4 + [C.c_void_p]),
    'mopo_sac_update': (C.c_int, [C.c_void_p, C.POINTER(SacBatch), C.c_void_p]),
    'mopo_sac_update_from_pools': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_uint64, C.c_int32,
                                             C.c_void_p, C.c_void_p]),
    'mopo_sac_export_policy': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'mopo_sac_opt_sizes': (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    'mopo_sac_get_opt_state': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_uint64), C.c_void_p]),
    'mopo_sac_set_opt_state': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int64, C.c_uint64, C.c_void_p]),
    'mopo_policy_create': (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    'mopo_policy_destroy': (C.c_int, [C.c_void_p]),
    'mopo_policy_set_params': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int, C.c_void_p]),
    'mopo_policy_act': (C.c_int, [C.c_void_p, C.POINTER(PolicyArgs), C.c_void_p]),
}

# developer probes: only in libmopo_b200_dbg.so (include/mopo_b200_debug.h; python -m mopo_b200.build --probes)
DEBUG_SYMBOLS = {
    'mopo_debug_umma_probe': (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'mopo_debug_tc_trace': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mopo_debug_tc3_trace': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    'mopo_debug_mma_interference': (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    'mopo_debug_pipe_bench': (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    'mopo_debug_mma_rate': (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    'mopo_debug_sac_trace': (C.c_int, [C.c_void_p, C.c_void_p]),
}

_lib = None
_dbg = None


def load():
    """Load the CUDA library; raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError('mopo_b200: %s is missing -- build it with `python -m mopo_b200.build` '
                          '(or __graft_entry__.build()); there is no CPU fallback' % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI is incomplete
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def load_probes():
    """The probe build of the library (all product symbols + mopo_debug_*): built on demand, never used by the product path."""
    global _dbg
    if _dbg is not None:
        return _dbg
    from . import build as _b
    path = _b.build(probes=True)
    lib = C.CDLL(path, mode=C.RTLD_LOCAL)
    for name, (res, args) in list(SYMBOLS.items()) + list(DEBUG_SYMBOLS.items()):
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _dbg = lib
    return lib


_EXC = {EINVAL: ValueError, ESTATE: RuntimeError, ECUDA: RuntimeError, EUNSUPPORTED: NotImplementedError}


def check(rc, lib=None):
    """Re-raise a C status as the exception type the reference would raise."""
    if rc != OK:
        msg = (lib or load()).mopo_last_error()
        raise _EXC.get(rc, RuntimeError)((msg or b'').decode() or 'mopo_b200 error %d' % rc)


def ptr(t):
    """Device (or host) address of a torch tensor / numpy array, None -> NULL."""
    if t is None:
        return None
    if hasattr(t, 'data_ptr'):
        return t.data_ptr()
    return t.ctypes.data


def stream_ptr(device):
    import torch
    return torch.cuda.current_stream(device).cuda_stream
