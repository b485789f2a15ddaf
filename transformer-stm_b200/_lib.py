"""ctypes binding of libcvt_b200.so (include/cvt_b200.h).  No CPU fallback: if the library is missing this raises."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libcvt_b200.so")
MAX_STAGES = 4
DATASET_SLOTS = 4
ABI_VERSION = 3

PROJ = {"dw_bn": 0, "avg": 1, "linear": 2, "linear_dense": 3}
HEAD = {"auto": 0, "cls_ln": 1, "ln_mean": 2, "cls": 3, "mean": 4, "gap": 4}
PREC = {"fp32": 0, "tf32": 1, "bf16": 2, "tf32x3": 3}
DROPOUT_OFF, DROPOUT_PHILOX, DROPOUT_MASKS = 0, 1, 2


class StageConfig(C.Structure):
    _fields_ = [("embed_dim", C.c_int32), ("patch_size", C.c_int32), ("stride", C.c_int32),
                ("num_heads", C.c_int32), ("qkv_method", C.c_int32), ("with_cls_token", C.c_int32)]


class Config(C.Structure):
    _fields_ = [("image_h", C.c_int32), ("image_w", C.c_int32), ("in_chans", C.c_int32), ("n_stages", C.c_int32),
                ("stage", StageConfig * MAX_STAGES),
                ("proc_dim", C.c_int32), ("proc_hidden", C.c_int32), ("num_classes", C.c_int32), ("mlp_ratio", C.c_int32),
                ("drop_rate", C.c_float), ("ln_eps", C.c_float), ("bn_eps", C.c_float), ("bn_momentum", C.c_float),
                ("precision", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32),
                ("head_kind", C.c_int32), ("skip_proj_qkv", C.c_int32), ("attn_drop_rate", C.c_float)]


class HostArrays(C.Structure):
    """cvt_host_arrays: the numpy arrays model.fit receives (float32 or float64), by pointer."""
    _fields_ = [("img", C.c_void_p), ("proc", C.c_void_p), ("y", C.c_void_p),
                ("img_f64", C.c_int32), ("proc_f64", C.c_int32), ("y_f64", C.c_int32), ("n", C.c_int64)]


class CvtError(RuntimeError):
    pass


_PROTOS = {
    "cvt_abi_version": (C.c_int, []),
    "cvt_default_config": (None, [C.POINTER(Config)]),
    "cvt_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "cvt_destroy": (None, [C.c_void_p]),
    "cvt_last_error": (C.c_char_p, []),
    "cvt_param_count": (C.c_int, [C.c_void_p]),
    "cvt_param_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int32), C.POINTER(C.c_int64),
                                 C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "cvt_param_index": (C.c_int, [C.c_void_p, C.c_char_p]),
    "cvt_set_param": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "cvt_get_param": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "cvt_get_grad": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "cvt_init_params": (C.c_int, [C.c_void_p, C.c_uint64]),
    "cvt_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "cvt_predict_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "cvt_loss_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "cvt_adam_step": (C.c_int, [C.c_void_p, C.c_float]),
    "cvt_train_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float]),
    "cvt_train_epoch_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "cvt_train_step_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float,
                                      C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "cvt_evaluate_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "cvt_dataset_alloc": (C.c_int, [C.c_void_p, C.c_int, C.c_int64]),
    "cvt_dataset_free": (C.c_int, [C.c_void_p, C.c_int]),
    "cvt_dataset_count": (C.c_int64, [C.c_void_p, C.c_int]),
    "cvt_dataset_write": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.POINTER(HostArrays)]),
    "cvt_dataset_write_u8": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "cvt_dataset_read": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cvt_train_epoch_stream": (C.c_int, [C.c_void_p, C.POINTER(HostArrays), C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_int,
                                         C.c_void_p, C.c_void_p]),
    "cvt_train_epoch_resident": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "cvt_predict_arrays": (C.c_int, [C.c_void_p, C.POINTER(HostArrays), C.c_void_p]),
    "cvt_evaluate_arrays": (C.c_int, [C.c_void_p, C.POINTER(HostArrays), C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "cvt_evaluate_resident": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "cvt_predict_resident": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "cvt_read_metrics": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "cvt_optimizer_reset": (C.c_int, [C.c_void_p]),
    "cvt_optimizer_step_count": (C.c_int64, [C.c_void_p]),
    "cvt_set_dropout": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64]),
    "cvt_set_dropout_mask": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64]),
    "cvt_get_tap": (C.c_int64, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64]),
    "cvt_backward_to_tap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64]),
    "cvt_comm_unique_id": (C.c_int, [C.c_void_p]),
    "cvt_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "cvt_comm_destroy": (C.c_int, [C.c_void_p]),
    "cvt_comm_peer_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cvt_comm_peer_open": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "cvt_comm_peer_close": (C.c_int, [C.c_void_p]),
    "cvt_gemm_dispatch_counts": (C.c_int, [C.POINTER(C.c_int64), C.c_int]),
    "cvt_grad_arena": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "cvt_param_arena": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "cvt_synchronize": (C.c_int, [C.c_void_p]),
    "cvt_stream": (C.c_void_p, [C.c_void_p]),
    "cvt_set_graphs": (C.c_int, [C.c_void_p, C.c_int]),
    "cvt_kernel_launches": (C.c_int64, [C.c_void_p]),
    "cvt_workspace_bytes": (C.c_int64, [C.c_void_p]),
    "cvt_debug_gemm": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cvt_debug_attention": (C.c_int, [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 9),
    "cvt_attention_backward_static": (C.c_int, [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 10),
    "cvt_profile_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "cvt_profile_group_launches": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
}

_lib = None


def exported_names():
    return sorted(_PROTOS)


def load():
    """Loads the shared library (building nothing: run __graft_entry__.build() first)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CvtError("libcvt_b200.so is not built (%s); run `python __graft_entry__.py` — there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.cvt_abi_version() != ABI_VERSION:
        raise CvtError("libcvt_b200.so has ABI %d, this binding expects %d: rebuild with `python __graft_entry__.py`" % (lib.cvt_abi_version(), ABI_VERSION))
    _lib = lib
    return lib


def check(rc, what=""):
    if rc is not None and rc < 0:
        msg = load().cvt_last_error()
        raise CvtError("%s failed (%d): %s" % (what or "cvt call", rc, msg.decode() if msg else "?"))
    return rc
