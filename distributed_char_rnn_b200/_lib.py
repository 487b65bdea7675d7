"""ctypes binding of ``libcrnn.so`` (the C-ABI declared in ``include/crnn.h``).

There is no fallback: if the library is missing this module raises on first use with
the build command, and the library itself refuses to create a handle without an
sm_100 GPU.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcrnn.so")

MODEL_IDS = {"rnn": 0, "gru": 1, "lstm": 2}
PREC_IDS = {"fp32": 0, "bf16": 1, "bf16x3": 2}
GRAD_TAIL = 2
NUM_PHASES = 12
CRNN_ERR_ARG, CRNN_ERR_CUDA, CRNN_ERR_UNSUP = -1, -2, -3        # include/crnn.h

c_i32, c_i64, c_f32, c_vp = ctypes.c_int32, ctypes.c_int64, ctypes.c_float, ctypes.c_void_p


class CrnnConfig(ctypes.Structure):
    _fields_ = [("model", c_i32), ("rnn_size", c_i32), ("num_layers", c_i32), ("seq_length", c_i32),
                ("batch_size", c_i32), ("vocab_size", c_i32), ("precision", c_i32), ("device", c_i32),
                ("input_keep_prob", c_f32), ("output_keep_prob", c_f32), ("dropout_seed", c_i32), ("reserved", c_i32)]


# name -> (restype, argtypes); must list every symbol include/crnn.h declares
SIGNATURES = {
    "crnn_create": (c_i32, [ctypes.POINTER(CrnnConfig), ctypes.POINTER(c_vp)]),
    "crnn_destroy": (None, [c_vp]),
    "crnn_last_error": (ctypes.c_char_p, []),
    "crnn_version": (ctypes.c_char_p, []),
    "crnn_flat_size": (c_i64, [c_vp]),
    "crnn_num_tensors": (c_i32, [c_vp]),
    "crnn_param_layout": (c_i32, [c_vp, c_i32, ctypes.c_char_p, c_i32, ctypes.POINTER(c_i64),
                                  ctypes.POINTER(c_i64), ctypes.POINTER(c_i64)]),
    "crnn_state_size": (c_i64, [c_vp]),
    "crnn_workspace_bytes": (c_i64, [c_vp]),
    "crnn_forward_backward": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "crnn_set_bucket_event": (c_i32, [c_vp, c_i32, c_vp]),
    "crnn_clip_adam": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_f32, c_f32, c_i64, c_f32, c_vp, c_vp, c_vp]),
    "crnn_clip_adam_noalloc": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_f32, c_f32, c_i64, c_f32, c_vp, c_vp]),
    "crnn_train_step": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_f32, c_f32, c_i64,
                                c_vp, c_vp]),
    "crnn_forward": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "crnn_launch_count": (ctypes.c_longlong, []),
    "crnn_profile_enable": (c_i32, [c_vp, c_i32]),
    "crnn_profile_read": (c_i32, [c_vp, ctypes.POINTER(ctypes.c_double), c_i32]),
    "crnn_profile_phase_name": (ctypes.c_char_p, [c_i32]),
    "crnn_profile_read_kernels": (c_i32, [c_vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_i64), c_i32]),
    "crnn_dropout_mask": (c_i32, [c_vp, c_i64, c_i32, c_vp, c_vp]),
    "crnn_embedding_gather": (c_i32, [c_vp, c_i32, c_i32, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "crnn_embedding_scatter_add": (c_i32, [c_vp, c_i32, c_i32, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "crnn_softmax_xent": (c_i32, [c_vp, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_vp]),
    "crnn_gemm_bf16": (c_i32, [c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp,
                               c_i32, c_vp]),
    "crnn_gemm_bf16x3": (c_i32, [c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_i64, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp,
                                 c_i32, c_vp]),
    "crnn_gemm_f32": (c_i32, [c_i32, c_i32, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp,
                              c_i32, c_vp]),
}

_lib = None


class CrnnError(RuntimeError):
    pass


def load():
    """Load ``libcrnn.so`` once and attach the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CrnnError("%s is missing - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise CrnnError("libcrnn: %s (code %d)" % (load().crnn_last_error().decode("utf-8", "replace"), rc))


def ptr(t):
    """Device (or host) address of a torch tensor / None."""
    return None if t is None else c_vp(t.data_ptr())
