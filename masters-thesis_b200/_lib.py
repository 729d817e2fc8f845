"""ctypes binding of libtnt.so (include/tnt.h).  There is no CPU fallback: if the shared library is
missing the import of this module raises, and tnt_create refuses to run without an sm_100 device."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtnt.so")


class TntConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("N", "R", "G", "U", "E", "A", "H", "V", "T", "B_max", "n_enc", "precision")] + \
               [(n, C.c_float) for n in ("dropout_input", "dropout_features", "dropout_text", "dropout_attn",
                                         "dropout_lstm", "dropout_out", "input_reg", "attn_reg", "lstm_reg",
                                         "output_reg", "lr", "beta1", "beta2", "adam_eps", "clipnorm")] + \
               [(n, C.c_int32) for n in ("emb_unaggregated_norm", "second_feature_dropout", "variant", "attn_loss")] + \
               [(n, C.c_float) for n in ("agc_clip", "agc_eps")]


class TntDenseConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("K", "E", "B_max")] + \
               [(n, C.c_float) for n in ("dropout_input", "dropout", "input_reg", "lr", "beta1", "beta2", "adam_eps",
                                         "clipnorm")]


class TntVarInfo(C.Structure):
    _fields_ = [("name", C.c_char * 64), ("offset", C.c_int64), ("rows", C.c_int64), ("cols", C.c_int64),
                ("l2", C.c_float), ("trainable", C.c_int32)]


class TntMetrics(C.Structure):
    _fields_ = [("ce_sum", C.c_double * 2), ("correct", C.c_double * 2), ("attn_mse", C.c_double * 2),
                ("l2", C.c_double), ("emb_sqnorm", C.c_double), ("rows", C.c_double * 2),
                ("attn_elem_mse", C.c_double)]


# every symbol include/tnt.h declares: (restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    "tnt_create": (C.c_int, [C.POINTER(TntConfig), _P, _P, C.POINTER(_P)]),
    "tnt_destroy": (C.c_int, [_P]),
    "tnt_last_error": (C.c_char_p, [_P]),
    "tnt_param_floats": (C.c_int64, [_P]),
    "tnt_state_floats": (C.c_int64, [_P]),
    "tnt_num_vars": (C.c_int, [_P]),
    "tnt_var_info_get": (C.c_int, [_P, C.c_int, C.POINTER(TntVarInfo)]),
    "tnt_bind_params": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "tnt_set_adam_step": (C.c_int, [_P, C.c_int64]),
    "tnt_get_adam_step": (C.c_int64, [_P]),
    "tnt_set_lr": (C.c_int, [_P, C.c_float]),
    "tnt_train_step": (C.c_int, [_P, _P, _P, _P, _P, _P, C.c_int, C.c_uint64, _P]),
    "tnt_eval_step": (C.c_int, [_P, _P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_uint64, _P]),
    "tnt_step": (C.c_int, [_P, _P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, _P]),
    "tnt_dp_unique_id": (C.c_int, [_P]),
    "tnt_dp_init": (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    "tnt_dp_world": (C.c_int, [_P]),
    "tnt_forward": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_uint64, _P, _P, _P]),
    "tnt_greedy": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, _P]),
    "tnt_sample": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_uint64, _P, _P, _P, _P, _P]),
    "tnt_beam_search": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, _P]),
    "tnt_check": (C.c_int, [_P, _P]),
    "tnt_kernel_variants": (C.c_char_p, [_P]),
    "tnt_set_attn_loss": (C.c_int, [_P, C.c_int]),
    "tnt_sam_perturb": (C.c_int, [_P, C.c_float, C.c_float, _P]),
    "tnt_sam_restore": (C.c_int, [_P, _P]),
    "tnt_read_metrics": (C.c_int, [_P, C.POINTER(TntMetrics), _P]),
    "tnt_onehot_to_ids": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "tnt_phase_prologue": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_uint64, _P]),
    "tnt_phase_encode": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_uint64, _P]),
    "tnt_phase_decode": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, _P]),
    "tnt_phase_encoder_backward": (C.c_int, [_P, _P, C.c_int, C.c_int, _P]),
    "tnt_phase_optimizer": (C.c_int, [_P, _P]),
    "tnt_phase_optimizer_part": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P]),
    "tnt_trace": (C.c_int, [_P, C.c_int]),
    "tnt_trace_read": (C.c_int, [_P, _P, C.c_int, _P, C.c_int]),
    "tnt_debug_encoder_trace": (None, [_P, _P]),
    "tnt_side_stream": (_P, [_P]),
    "tnt_stream_wait_event": (C.c_int, [_P, C.c_int, _P]),
    "tnt_side_join": (C.c_int, [_P, _P]),
    "tnt_eval_l2": (C.c_int, [_P, _P]),
    "tnt_buffer": (C.c_int, [_P, C.c_char_p, C.POINTER(_P), C.POINTER(C.c_int64)]),
    "tnt_grad_bucket": (C.c_int, [_P, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "tnt_launch_count": (C.c_int64, [_P]),
    # dense voxel encoder of NIC.py (SURVEY 8a A1'')
    "tnt_dense_create": (C.c_int, [C.POINTER(TntDenseConfig), C.POINTER(_P)]),
    "tnt_dense_destroy": (C.c_int, [_P]),
    "tnt_dense_last_error": (C.c_char_p, [_P]),
    "tnt_dense_param_floats": (C.c_int64, [_P]),
    "tnt_dense_launch_count": (C.c_int64, [_P]),
    "tnt_dense_bind_params": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "tnt_dense_set_adam_step": (C.c_int, [_P, C.c_int64]),
    "tnt_dense_set_lr": (C.c_int, [_P, C.c_float]),
    "tnt_dense_forward": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_uint64, _P, _P]),
    "tnt_dense_forward_noalloc": (C.c_int, [_P, _P, C.c_int, _P, _P]),
    "tnt_dense_backward": (C.c_int, [_P, _P, _P]),
    "tnt_dense_apply_gradients": (C.c_int, [_P, _P]),
    # batch assembly from HBM-resident tables
    "tnt_assemble_betas": (C.c_int, [_P, C.c_int64, C.c_int64, _P, C.c_int, _P, _P, _P]),
    "tnt_assemble_captions": (C.c_int, [_P, C.c_int64, C.c_int, _P, C.c_int, _P, _P, _P, C.c_int, _P, _P]),
    "tnt_data_last_error": (C.c_char_p, []),
    "tnt_data_launch_count": (C.c_int64, []),
}


def load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback for the Think-and-Tell hot path)")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


class TntError(RuntimeError):
    pass


def check(lib, ctx, rc):
    if rc != 0:
        msg = lib.tnt_last_error(ctx)
        raise TntError(f"libtnt error {rc}: {msg.decode() if msg else '?'}")
