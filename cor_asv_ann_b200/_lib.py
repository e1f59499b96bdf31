"""ctypes binding of libcor_b200.so (include/cor_b200.h).

There is no CPU fallback: importing this module without the built library, or
calling into it without a CUDA device, raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcor_b200.so")


class CorConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "width", "depth", "voc_size", "residual_connections", "deep_bidirectional_encoder",
        "bridge_dense", "lm_predict", "window_width", "eos_index", "gemm_mode")]


class CorLines(C.Structure):
    _fields_ = [("n_rows", C.c_int32), ("t_max", C.c_int32), ("k_alt", C.c_int32),
                ("row_T", C.c_void_p), ("idx", C.c_void_p), ("val", C.c_void_p)]


class CorBeamParams(C.Structure):
    _fields_ = [("batch_size", C.c_int32), ("beam_width_in", C.c_int32), ("beam_width_out", C.c_int32),
                ("n_best", C.c_int32), ("beam_threshold_in", C.c_double), ("rejection_threshold", C.c_double)]


class CorTrainBatch(C.Structure):
    _fields_ = [("n_rows", C.c_int32), ("t_src", C.c_int32), ("k_alt", C.c_int32), ("t_dec", C.c_int32),
                ("src_idx", C.c_void_p), ("src_val", C.c_void_p), ("dec_in", C.c_void_p), ("dec_out", C.c_void_p),
                ("mask_enc", C.c_void_p), ("mask_dec", C.c_void_p), ("mask_cell", C.c_void_p)]


# name -> (restype, argtypes); every symbol include/cor_b200.h declares
PROTOTYPES = {
    "cor_last_error": (C.c_char_p, []),
    "cor_abi_version": (C.c_int, []),
    "cor_model_create": (C.c_int, [C.POINTER(CorConfig), C.c_int, C.POINTER(C.c_void_p)]),
    "cor_model_set_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_void_p, C.POINTER(C.c_int64), C.c_int]),
    "cor_model_finalize": (C.c_int, [C.c_void_p]),
    "cor_model_destroy": (None, [C.c_void_p]),
    "cor_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "cor_encode": (C.c_int, [C.c_void_p, C.POINTER(CorLines), C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cor_set_encoding": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "cor_decode_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p] + [C.c_void_p] * 9 + [C.c_void_p]),
    "cor_decode_greedy_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t] + [C.c_void_p] * 6 + [C.c_void_p]),
    "cor_correct_greedy_host": (C.c_int, [C.c_void_p, C.POINTER(CorLines), C.c_void_p, C.c_size_t] + [C.c_void_p] * 6 + [C.c_void_p]),
    "cor_beam_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(CorBeamParams), C.c_size_t, C.POINTER(C.c_size_t)]),
    "cor_correct_beam_host": (C.c_int, [C.c_void_p, C.POINTER(CorLines), C.POINTER(CorBeamParams), C.c_void_p, C.c_size_t] + [C.c_void_p] * 6 + [C.c_void_p]),
    "cor_decode_beam_batch": (C.c_int, [C.c_void_p, C.POINTER(CorLines), C.POINTER(CorBeamParams), C.c_void_p, C.c_size_t] + [C.c_void_p] * 6 + [C.c_void_p]),
    "cor_alignment_paths": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                      C.c_double, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "cor_trainer_create": (C.c_int, [C.POINTER(CorConfig), C.c_int, C.POINTER(C.c_void_p)]),
    "cor_trainer_destroy": (None, [C.c_void_p]),
    "cor_trainer_param_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "cor_trainer_slot": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "cor_trainer_bind": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cor_trainer_reset_optimizer": (C.c_int, [C.c_void_p]),
    "cor_trainer_launch_count": (C.c_int64, [C.c_void_p, C.c_int]),
    "cor_train_forward_backward": (C.c_int, [C.c_void_p, C.POINTER(CorTrainBatch), C.POINTER(C.c_double), C.c_void_p]),
    "cor_train_apply": (C.c_int, [C.c_void_p, C.c_double, C.c_int, C.c_float, C.c_float, C.POINTER(C.c_double), C.c_void_p]),
    "cor_launch_count": (C.c_int64, [C.c_void_p, C.c_int]),
    "cor_host_launch_count": (C.c_int64, [C.c_void_p, C.c_int]),
    "cor_debug_counters": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.c_int]),
    "cor_debug_gemm": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cor_gemm_noalloc": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cor_gemm_noalloc_max_bytes": (C.c_size_t, []),
    "cor_prof_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "cor_prof_read": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_double)]),
}

_lib = None


def load():
    """dlopen the product library and bind every declared entry point."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s is missing: build it with `make` (or __graft_entry__.build()); "
                           "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)       # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class CorError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise CorError(load().cor_last_error().decode("utf-8", "replace"))
