"""ctypes binding of libsmcphys.so (include/smcphys.h).  Loading fails loudly: there is no CPU path."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# SMCPHYS_SO points at an alternative build of the same library (kernel tuning experiments)
SO = os.environ.get("SMCPHYS_SO") or os.path.join(HERE, "libsmcphys.so")

SPHERE, CUBE = 0, 1


class SmcphysError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libsmcphys error %d: %s" % (code, msg))
        self.code = code


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("device", C.c_int32), ("shape", C.c_int32), ("substeps", C.c_int32),
        ("num_particles", C.c_int64), ("global_particles", C.c_int64), ("particle_offset", C.c_int64),
        ("max_timesteps", C.c_int32), ("rejuvenation_moves", C.c_int32),
        ("observation_noise", C.c_double), ("transition_noise", C.c_double), ("proposal_sigma", C.c_double),
        ("ceiling_half", C.c_double), ("ess_num", C.c_int64), ("ess_den", C.c_int64), ("seed", C.c_uint64),
        ("resampler", C.c_int32), ("pad_", C.c_int32),
    ]


class StepStats(C.Structure):
    _fields_ = [("ess", C.c_double), ("lse", C.c_double), ("accept_rate", C.c_double), ("resampled", C.c_int32),
                ("pad_", C.c_int32)]


_dp, _fp = C.POINTER(C.c_double), C.POINTER(C.c_float)
_ip, _up = C.POINTER(C.c_int64), C.POINTER(C.c_uint64)
_H = C.c_void_p

# name -> (restype, argtypes); every symbol include/smcphys.h declares
SIGNATURES = {
    "smcphys_config_default": (C.c_int, [C.POINTER(Config)]),
    "smcphys_last_error": (C.c_char_p, []),
    "smcphys_device_count": (C.c_int, []),
    "smcphys_launch_count": (C.c_int64, []),
    "smcphys_create": (C.c_int, [C.POINTER(Config), C.POINTER(_H)]),
    "smcphys_create_multi": (C.c_int, [C.POINTER(Config), C.c_int32, C.POINTER(C.c_int32), C.POINTER(_H)]),
    "smcphys_comm_unique_id": (C.c_int, [C.c_char_p]),
    "smcphys_comm_init_rank": (C.c_int, [_H, C.c_char_p, C.c_int32, C.c_int32]),
    "smcphys_total_particles": (C.c_int64, [_H]),
    "smcphys_steps_done": (C.c_int32, [_H]),
    "smcphys_destroy": (C.c_int, [_H]),
    "smcphys_set_stream": (C.c_int, [_H, C.c_void_p]),
    "smcphys_set_trial": (C.c_int, [_H, _dp, _dp, _dp, C.c_int32]),
    "smcphys_run_smc": (C.c_int, [_H, C.c_int32, C.POINTER(StepStats)]),
    "smcphys_get_particles": (C.c_int, [_H, _dp, _dp, _dp, _fp]),
    "smcphys_get_states": (C.c_int, [_H, _fp]),
    "smcphys_get_log_ml": (C.c_int, [_H, _dp]),
    "smcphys_predict": (C.c_int, [_H, C.c_int32, _fp, _fp]),
    "smcphys_pf_init": (C.c_int, [_H]),
    "smcphys_pf_extend": (C.c_int, [_H, C.c_int32, C.c_void_p]),
    "smcphys_pf_weight_sums": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "smcphys_ess_below": (C.c_int, [C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_int64]),
    "smcphys_ess_value": (C.c_double, [C.c_uint64, C.c_uint64]),
    "smcphys_lse_value": (C.c_double, [C.c_double, C.c_uint64]),
    "smcphys_resample_offset": (C.c_double, [C.c_uint64, C.c_int32]),
    "smcphys_pf_offspring_range": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int64, C.c_double, _ip, _ip]),
    "smcphys_pf_payload_words": (C.c_int64, [_H, C.c_int32]),
    "smcphys_pf_pack": (C.c_int, [_H, C.c_int32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double, C.c_int64, C.c_int64,
                                  C.c_int64, C.c_void_p]),
    "smcphys_pf_gather_local": (C.c_int, [_H, C.c_int32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double, C.c_int64,
                                          C.c_int64, C.c_int64]),
    "smcphys_pf_unpack": (C.c_int, [_H, C.c_int32, C.c_void_p, C.c_int32, _ip, _ip, _ip]),
    "smcphys_pf_resample_local": (C.c_int, [_H, C.c_int32]),
    "smcphys_pf_rejuvenate": (C.c_int, [_H, C.c_int32, C.c_int32, C.c_void_p]),
    "smcphys_pf_rejuvenate_from": (C.c_int, [_H, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "smcphys_synchronize": (C.c_int, [_H]),
    "smcphys_sim_step": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_int64, _fp, _fp]),
    "smcphys_simulate": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double, _dp, _dp, _fp, _fp,
                                   C.c_int64, C.c_int32, _fp, _dp, _fp]),
    "smcphys_resample_systematic": (C.c_int, [C.c_int32, _dp, C.c_int64, C.c_double, _ip]),
    "smcphys_resample_multinomial": (C.c_int, [C.c_int32, _dp, C.c_int64, C.c_uint64, C.c_int32, _ip]),
    "smcphys_resample_systematic_noalloc": (C.c_int, [C.c_int32, C.c_void_p, C.c_int64, C.c_double, C.c_void_p, C.c_void_p]),
    "smcphys_weight_stats": (C.c_int, [C.c_int32, _dp, C.c_int64, _dp, _up]),
    "smcphys_set_profiling": (C.c_int, [_H, C.c_int32]),
    "smcphys_kernel_time": (C.c_int, [_H, C.c_int32, _dp, _ip]),
    "smcphys_fp32_peak": (C.c_int, [C.c_int32, _dp]),
}

_LIB = None


def lib():
    """The loaded library.  Raises if it has not been built: the product has no other code path."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO):
            raise RuntimeError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(needs nvcc; libsmcphys has no CPU fallback)" % SO)
        L = C.CDLL(SO)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise SmcphysError(rc, lib().smcphys_last_error().decode("utf-8", "replace"))
    return rc
