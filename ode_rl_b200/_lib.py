"""ctypes binding of libodex.so (the C-ABI declared in include/odex.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make``.  There is NO fallback: if the
shared object is missing, or a symbol the header declares is not exported, importing the product
modules on a CUDA box fails loudly.  (On a CPU-only box the library still loads -- it links the CUDA
runtime statically -- so the symbol table can be checked without a GPU; no compute call is made.)
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libodex.so")

ODEX_VERSION = 200
ACT_NONE, ACT_RELU, ACT_TANH = 0, 1, 2
CONV_AUTO, CONV_SIMT, CONV_TC = 0, 1, 2
STATE_HEAD_BYTES = 256

c_float_p = C.POINTER(C.c_float)
c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class ConvDesc(C.Structure):
    """struct odex_conv_desc (include/odex.h)."""

    _fields_ = [
        ("x0", C.c_void_p), ("ld0", C.c_int), ("c0", C.c_int),
        ("x1", C.c_void_p), ("ld1", C.c_int), ("c1", C.c_int),
        ("B", C.c_int), ("H", C.c_int), ("W", C.c_int), ("R", C.c_int), ("N", C.c_int),
        ("w_simt", C.c_void_p),
        ("w_tc", C.c_void_p),
        ("bias", C.c_void_p),
        ("act", C.c_int),
        ("dact", C.c_void_p), ("ldd", C.c_int), ("dact_kind", C.c_int),
        ("res", C.c_void_p), ("ldr", C.c_int),
        ("alpha", C.c_float), ("alpha_dev", C.c_void_p),
        ("out", C.c_void_p), ("ldo", C.c_int),
        ("skip", C.c_void_p),
        ("impl", C.c_int),
    ]


NET_MAX_LAYERS = 6


class ConvNetDesc(C.Structure):
    """struct odex_convnet_desc (include/odex.h)."""

    _fields_ = [
        ("x", C.c_void_p), ("ldx", C.c_int),
        ("B", C.c_int), ("H", C.c_int), ("W", C.c_int), ("C", C.c_int), ("L", C.c_int),
        ("w_tc", C.c_void_p * NET_MAX_LAYERS),
        ("bias", C.c_void_p * NET_MAX_LAYERS),
        ("act", C.c_int * NET_MAX_LAYERS),
        ("aux", C.c_void_p * NET_MAX_LAYERS), ("ld_aux", C.c_int * NET_MAX_LAYERS), ("aux_kind", C.c_int * NET_MAX_LAYERS),
        ("out", C.c_void_p * NET_MAX_LAYERS), ("ld_out", C.c_int * NET_MAX_LAYERS),
        ("alpha", C.c_float), ("alpha_dev", C.c_void_p),
        ("skip", C.c_void_p),
        ("split", C.c_int),
        ("amax_out", C.c_void_p * NET_MAX_LAYERS),
    ]


C16_PLAIN, C16_GRU_GATES, C16_GRU_OUT = 0, 1, 2


class Conv16Desc(C.Structure):
    """struct odex_conv16_desc (include/odex.h)."""

    _fields_ = [
        ("x0", C.c_void_p), ("ld0", C.c_int),
        ("x1", C.c_void_p), ("ld1", C.c_int),
        ("B", C.c_int), ("R", C.c_int), ("N", C.c_int),
        ("w_f16", C.c_void_p),
        ("bias", C.c_void_p),
        ("mode", C.c_int),
        ("act", C.c_int), ("aux", C.c_void_p), ("ld_aux", C.c_int), ("aux_kind", C.c_int), ("alpha", C.c_float),
        ("alpha_dev", C.c_void_p),
        ("out", C.c_void_p), ("ldo", C.c_int),
        ("gamma", C.c_void_p), ("beta", C.c_void_p), ("eps", C.c_float),
        ("h", C.c_void_p), ("ldh", C.c_int), ("z_in", C.c_void_p),
        ("z", C.c_void_p), ("rh", C.c_void_p), ("hnew", C.c_void_p), ("ld_hnew", C.c_int), ("cand", C.c_void_p),
        ("g_raw", C.c_void_p), ("stats", C.c_void_p),
        ("skip", C.c_void_p),
        ("out_hi", C.c_void_p), ("ld_out_hi", C.c_int), ("out_split", C.c_int),
        ("amax_out", C.c_void_p),
    ]


class DopriState(C.Structure):
    """struct odex_dopri_state (include/odex.h) -- the host reads a copy of it back once per forward."""

    _fields_ = [
        ("t0", C.c_double), ("t1", C.c_double), ("dt", C.c_double), ("dt_try", C.c_double),
        ("rtol", C.c_double), ("atol", C.c_double),
        ("red", C.c_double * 4),
        ("ratio", C.c_double),
        ("h0", C.c_double), ("d0", C.c_double), ("d1", C.c_double), ("d2", C.c_double),
        ("accept", C.c_int), ("done", C.c_int), ("adv_seq", C.c_int), ("jnext", C.c_int),
        ("jlo", C.c_int), ("jhi", C.c_int), ("n_out", C.c_int),
        ("n_accept", C.c_int), ("n_reject", C.c_int), ("n_feval", C.c_int), ("n_attempt", C.c_int),
        ("status", C.c_int), ("max_attempts", C.c_int), ("save_cap", C.c_int),
        ("dt_acc", C.c_double),
    ]


# byte offsets of fields the host addresses on the device (checked against the struct above)
OFF_DT = DopriState.dt.offset
OFF_DT_TRY = DopriState.dt_try.offset
OFF_RED = DopriState.red.offset
OFF_H0 = DopriState.h0.offset
OFF_DONE = DopriState.done.offset

# name -> (restype, argtypes); every symbol include/odex.h declares
_vp, _i, _f, _d, _ll, _u32, _sz = C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_longlong, C.c_uint32, C.c_size_t
SIGNATURES = {
    "odex_version": (_i, []),
    "odex_last_error": (C.c_char_p, []),
    "odex_launch_count": (_ll, []),
    "odex_packed_weight_floats": (_sz, [_i, _i, _i, _i]),
    "odex_pack_weights": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp]),
    "odex_pack_weights_f16_pair": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp]),
    "odex_conv2d": (_i, [C.POINTER(ConvDesc), _vp]),
    "odex_conv2d_repeat": (_i, [C.POINTER(ConvDesc), _i, _vp]),
    "odex_convnet": (_i, [C.POINTER(ConvNetDesc), _vp]),
    "odex_convnet_supported": (_i, [C.POINTER(ConvNetDesc)]),
    "odex_conv16": (_i, [C.POINTER(Conv16Desc), _vp]),
    "odex_conv16_supported": (_i, [C.POINTER(Conv16Desc)]),
    "odex_decode_tail_fwd": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _f, _vp, _vp]),
    "odex_decode_tail_bwd": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _f, _vp, _vp, _vp, _vp]),
    "odex_encode_stem_fwd": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _f, _vp, _vp]),
    "odex_encode_stem_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp]),
    "odex_bn_scratch_floats": (_ll, [_ll, _i]),
    "odex_bn_relu_fwd": (_i, [_vp, _ll, _i, _vp, _vp, _f, _i, _vp, _vp, _vp, _ll, _vp]),
    "odex_bn_static_scratch_floats": (_ll, []),
    "odex_bn_relu_fwd_static": (_i, [_vp, _ll, _i, _vp, _vp, _f, _i, _vp, _vp, _vp]),
    "odex_bn_relu_bwd": (_i, [_vp, _vp, _vp, _ll, _i, _vp, _vp, _i, _vp, _vp, _vp, _ll, _vp]),
    "odex_warp_compose_fwd": (_i, [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "odex_upsample2x_fwd": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "odex_upsample2x_bwd": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "odex_conv2d_wgrad": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "odex_conv2d_wgrad_packed": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "odex_wgrad_unpack": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "odex_conv2d_wgrad_f16": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "odex_conv2d_wgrad_f16_scaled": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "odex_gru_gates_fwd": (_i, [_vp, _vp, _i, _vp, _vp, _f, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "odex_gru_out_fwd": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _f, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "odex_gru_out_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "odex_gru_gates_bwd": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp]),
    "odex_lincomb": (_i, [_vp, _i, C.POINTER(_vp), c_double_p, _u32, _vp, _ll, _vp, _vp]),
    "odex_act_bwd": (_i, [_vp, _vp, _vp, _i, _ll, _vp]),
    "odex_dopri_state_bytes": (_sz, []),
    "odex_dopri_init": (_i, [_vp, _vp, _i, _d, _d, _i, _i, _vp]),
    "odex_dopri_initnorm": (_i, [_vp, _i, _vp, _vp, _vp, _ll, _i, _vp]),
    "odex_dopri_initctrl": (_i, [_vp, _i, _vp]),
    "odex_dopri_errctrl": (_i, [_vp, _vp, _vp, _vp, C.POINTER(_vp), _ll, _vp, _i, _i, _vp]),
    "odex_dopri_errnorm": (_i, [_vp, _vp, _vp, C.POINTER(_vp), _ll, _vp]),
    "odex_dopri_ctrl": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "odex_dopri_advance": (_i, [_vp, _vp, _vp, _vp, C.POINTER(_vp), _vp, _ll, _vp, _vp, _i, _i, _vp]),
    "odex_nan_flag": (_i, [_vp, _ll, _vp, _i, _vp]),
    "odex_nchw_to_nhwc": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "odex_nhwc_to_nchw": (_i, [_vp, _vp, _i, _i, _i, _vp]),
}


class OdexError(RuntimeError):
    pass


_lib = None


def lib() -> C.CDLL:
    """Load libodex.so once; raise (never fall back) when it is missing or incomplete."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise OdexError(
            "libodex.so not found at %s -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make`). There is no CPU / eager fallback for the ODE-RL hot path." % LIB_PATH)
    l = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(l, name)
        except AttributeError as e:  # pragma: no cover - build problem
            raise OdexError("libodex.so does not export %s (stale build?)" % name) from e
        fn.restype = res
        fn.argtypes = args
    v = l.odex_version()
    if v != ODEX_VERSION:
        raise OdexError("libodex.so version %d != binding version %d" % (v, ODEX_VERSION))
    _lib = l
    return l


def check(rc: int, what: str = "") -> None:
    """Turn the ABI's int status into an exception (0 ok, <0 CUDA error, >0 ODEX_E_*)."""
    if rc != 0:
        msg = lib().odex_last_error().decode("utf-8", "replace")
        raise OdexError("%s failed with status %d: %s" % (what or "odex call", rc, msg))


def launch_count() -> int:
    return int(lib().odex_launch_count())
