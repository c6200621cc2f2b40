"""ctypes binding of libb200fit.so (C ABI declared in include/b200fit.h).

There is no CPU fallback: importing this module fails loudly when the CUDA
library has not been built (run `python -c "import __graft_entry__ as g; g.build()"`
or `make -C orange3-spark_b200/csrc`).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2F_LIB_PATH: dev knob (kernel experiments built next to the product library), never a fallback
LIB_PATH = os.environ.get("B2F_LIB_PATH") or os.path.join(_HERE, "libb200fit.so")

# dtype codes (include/b200fit.h)
F32, F64, I32, I64, U8 = 0, 1, 2, 3, 4
KM_AUTO, KM_GENERIC, KM_TC3, KM_TC1, KM_TCB, KM_TCH = 0, 1, 2, 3, 4, 5

_vp, _i32, _i64, _u64, _f32, _f64, _sz = (C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float, C.c_double,
                                          C.c_size_t)

# name -> (restype, argtypes); every symbol include/b200fit.h declares
PROTOTYPES = {
    "b2f_version": (C.c_int, []),
    "b2f_last_error": (C.c_char_p, []),
    "b2f_synth_mixture": (C.c_int, [_vp, _i64, _i64, _i32, _i64, _u64, _i32, _f32, _vp, _vp]),
    "b2f_synth_logistic": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i64, _u64, _vp, _vp, _vp, _i32, _vp]),
    "b2f_pack_columns": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _i64, _vp, _vp]),
    "b2f_colstats": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _vp, _vp]),
    "b2f_scale_columns": (C.c_int, [_vp, _vp, _i64, _i32, _i64, _vp, _vp, _vp]),
    "b2f_silhouette_stats": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _i32, _vp, _vp, _vp]),
    "b2f_silhouette_scratch_bytes": (_sz, [_i32, _i32]),
    "b2f_silhouette_score": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _i32, _vp, _vp, _vp, _vp]),
    "b2f_label_hist": (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp]),
    "b2f_kmeans_prep_bytes": (_sz, [_i32, _i32]),
    "b2f_kmeans_acc_bytes": (_sz, [_i32, _i32]),
    "b2f_kmeans_payload_len": (_i64, [_i32, _i32]),
    "b2f_kmeans_prepare": (C.c_int, [_vp, _i32, _i32, _vp, _vp]),
    "b2f_kmeans_prepare_x": (C.c_int, [_vp, _i32, _i32, _f64, _vp, _vp]),
    "b2f_row_norm_max": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _vp]),
    "b2f_kmeans_acc_zero": (C.c_int, [_vp, _i32, _i32, _vp]),
    "b2f_kmeans_step": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _i32, _vp, _vp, _i32, _vp]),
    "b2f_kmeans_ws_bytes": (_sz, [_i64, _i32, _i32]),
    "b2f_kmeans_step_ws": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _i32, _vp, _vp, _vp, _sz, _i32, _vp]),
    "b2f_kmeans_scaled_supported": (C.c_int, [_i32, _i32, _i64]),
    "b2f_kmeans_step_scaled": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2f_kmeans_predict_ws": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _vp, _i32, _vp, _vp, _vp, _sz, _i32, _vp]),
    "b2f_kmeans_fold": (C.c_int, [_vp, _i32, _i32, _vp]),
    "b2f_kmeans_finalize": (C.c_int, [_vp, _i32, _i32, _f64, _vp, _vp, _vp, _vp]),
    "b2f_kmeans_predict": (C.c_int, [_vp, _i64, _i32, _i64, _vp, _i32, _vp, _vp, _i32, _vp]),
    "b2f_lr_scratch_bytes": (_sz, [_i32, _i32]),
    "b2f_lr_binary_lossgrad": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _i64, _vp, _vp, _vp, _vp, _vp]),
    "b2f_lr_binary_lossgrad_noalloc": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _i64, _vp, _vp, _vp, _vp]),
    "b2f_lr_multinomial_lossgrad": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _i64, _i32, _vp, _vp, _vp, _vp, _vp]),
    "b2f_lr_predict": (C.c_int, [_vp, _i64, _i32, _i64, _i32, _vp, _vp, _f64, _vp, _vp, _vp, _vp]),
}


class B200FitError(RuntimeError):
    pass


_lib = None


def lib():
    """Loads the shared library once.  Raises (never falls back) if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200FitError(
                "libb200fit.so not found at %s -- build it first (make -C orange3-spark_b200/csrc); "
                "b200fit has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().b2f_last_error()
        raise B200FitError("%s failed (%d): %s" % (what or "b200fit call", rc, msg.decode() if msg else "?"))


def ptr(t):
    """Device/host pointer of a torch tensor (or None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda():
    import torch

    if not torch.cuda.is_available():
        raise B200FitError("b200fit needs a CUDA device (B200, sm_100a); there is no CPU fallback")
