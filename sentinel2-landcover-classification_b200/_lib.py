"""ctypes binding of libs2lc.so (the C ABI declared in include/s2lc.h).

The library is the only compute backend: if it cannot be loaded this module raises — there is no
CPU or PyTorch fallback for the hot path.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libs2lc.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_SILU, ACT_RELU = 0, 1, 2
CONV_1X1, CONV_3X3, CONVT_FWD, CONVT_DGRAD = 0, 1, 2, 3
STATS_NONE, STATS_MOMENTS, STATS_BNBWD = 0, 1, 2
LOSS_FOCAL, LOSS_CE = 0, 1

vp = C.c_void_p
i32, i64, u64, f32, f64 = C.c_int32, C.c_int64, C.c_uint64, C.c_float, C.c_double


class Lazy(C.Structure):
    _fields_ = [("raw", vp), ("ld", i64), ("scale", vp), ("shift", vp), ("gate", vp), ("act", i32), ("_pad", i32)]


class LazyGrad(C.Structure):
    _fields_ = [("dy", vp), ("ld_dy", i64), ("z", vp), ("ld_z", i64), ("p", vp), ("q", vp), ("t", vp),
                ("rowscale", vp)]


class ConvParams(C.Structure):
    _fields_ = [("mode", i32), ("dtype", i32), ("N", i32), ("H", i32), ("W", i32), ("C", i32), ("Cpad", i32),
                ("Kpad", i32), ("Nout", i32), ("Npad", i32), ("_pad0", i32), ("a_is_grad", i32),
                ("a_fwd", C.POINTER(Lazy)), ("a_bwd", C.POINTER(LazyGrad)), ("w_packed", vp), ("bias", vp),
                ("out", vp), ("ld_out", i64), ("res1", vp), ("ld_res1", i64), ("res2", vp), ("ld_res2", i64),
                ("prev", C.POINTER(Lazy)), ("stats", vp), ("stats_kind", i32), ("_pad", i32),
                ("stats_rowscale", vp)]


class PackEntry(C.Structure):
    _fields_ = [("start", i64), ("numel", i64), ("dst", vp), ("stride", i64 * 4), ("role", i32 * 4), ("N2", i32),
                ("Tt", i32), ("Cpad", i32), ("Kpad", i32), ("flip", i32), ("dtype", i32)]


class WgradParams(C.Structure):
    _fields_ = [("mode", i32), ("dtype", i32), ("N", i32), ("H", i32), ("W", i32), ("Cin", i32), ("Cout", i32),
                ("_pad", i32), ("a", C.POINTER(Lazy)), ("g", C.POINTER(LazyGrad)), ("dw", vp), ("dbias", vp)]


_PROTOS = {
    "s2lc_version": ([], i32),
    "s2lc_last_error": ([], C.c_char_p),
    "s2lc_arch_check": ([i32], i32),
    "s2lc_memset_zero": ([vp, i64, vp], i32),
    "s2lc_weight_pack": ([vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i64, i64, i64, i64, i32, vp], i32),
    "s2lc_stem_conv_fwd": ([vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_stem_im2col": ([vp, vp, i64, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_stem_conv_wgrad": ([vp, C.POINTER(LazyGrad), vp, i32, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_bn_finalize": ([vp, vp, vp, f64, f32, f32, vp, vp, vp, vp, vp, vp, i32, vp], i32),
    "s2lc_bn_eval_affine": ([vp, vp, vp, vp, f32, vp, vp, i32, vp], i32),
    "s2lc_bn_apply": ([C.POINTER(Lazy), vp, i64, vp, vp, i64, i64, i32, i32, i32, vp], i32),
    "s2lc_bn_bwd_reduce": ([vp, i64, vp, i64, vp, vp, i64, i32, i32, i32, vp], i32),
    "s2lc_bn_bwd_finalize": ([vp, vp, vp, vp, f64, vp, vp, vp, vp, vp, i32, vp], i32),
    "s2lc_dwconv_fwd": ([C.POINTER(Lazy), vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_dwconv_bwd_data": ([C.POINTER(LazyGrad), vp, C.POINTER(Lazy), vp, i64, vp, i64, vp, i32, i32, i32, i32,
                              i32, i32, i32, vp], i32),
    "s2lc_dwconv_bwd_weight": ([C.POINTER(LazyGrad), C.POINTER(Lazy), vp, i32, i32, i32, i32, i32, i32, i32, vp],
                               i32),
    "s2lc_dwconv_bwd": ([C.POINTER(LazyGrad), vp, C.POINTER(Lazy), i32, vp, i64, vp, i64, vp, vp, i32, i32, i32, i32,
                         i32, i32, i32, vp], i32),
    "s2lc_dwconv_fwd_tile": ([C.POINTER(Lazy), vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_prepare_batch": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, vp], i32),
    "s2lc_argmax_confusion": ([vp, i64, vp, vp, vp, i64, i32, i32, vp], i32),
    "s2lc_lazygrad_apply": ([C.POINTER(LazyGrad), vp, i64, i64, i32, i32, i32, vp], i32),
    "s2lc_gather_tiles": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, vp], i32),
    "s2lc_argmax_mosaic": ([vp, i64, vp, vp, i64, i32, i32, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_se_pool": ([C.POINTER(Lazy), vp, i32, i32, i32, i32, vp], i32),
    "s2lc_se_fc_fwd": ([vp, vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
    "s2lc_se_fc_fwd_noalloc": ([vp, vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
    "s2lc_se_bwd_reduce": ([vp, i64, C.POINTER(Lazy), vp, i32, i32, i32, i32, vp], i32),
    "s2lc_se_fc_bwd": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
    "s2lc_se_bwd_apply": ([vp, i64, C.POINTER(Lazy), vp, vp, i32, i32, i32, i32, vp], i32),
    "s2lc_conv_fwd": ([C.POINTER(ConvParams), vp], i32),
    "s2lc_conv_wgrad": ([C.POINTER(WgradParams), vp], i32),
    "s2lc_outconv_fwd": ([C.POINTER(Lazy), vp, vp, vp, i64, i64, i32, i32, i32, vp], i32),
    "s2lc_outconv_bwd": ([C.POINTER(Lazy), vp, i64, vp, vp, vp, vp, i64, vp, i64, i32, i32, i32, vp], i32),
    "s2lc_copy_slice": ([vp, i64, vp, i64, i64, i32, i32, vp], i32),
    "s2lc_add_slice": ([vp, i64, vp, i64, vp, i64, i64, i32, i32, vp], i32),
    "s2lc_nchw_to_slice": ([vp, vp, i64, i32, i32, i32, i32, i32, i32, vp], i32),
    "s2lc_dropconnect_masks": ([vp, i32, i32, f32, u64, u64, vp], i32),
    "s2lc_loss_fwd_bwd": ([vp, i64, vp, vp, f32, f32, i64, i32, i32, vp, vp, vp, i64, i64, i32, i32, vp], i32),
    "s2lc_scale_inplace": ([vp, i64, vp, i64, i32, i32, vp], i32),
    "s2lc_adam_step": ([vp, vp, vp, vp, i64, f64, f64, f64, f64, f64, i64, f32, vp], i32),
    "s2lc_adam_pack_step": ([vp, vp, vp, vp, i64, f64, f64, f64, f64, f64, i64, f32, vp, i32, vp], i32),
    "s2lc_simt_fallbacks": ([i32], i64),
    "s2lc_tma_launches": ([i32], i64),
    "s2lc_num_sms": ([], i32),
    "s2lc_set_debug_flags": ([C.c_uint32], C.c_uint32),
}

EXPORTED_SYMBOLS = tuple(_PROTOS)

_lib = None


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load libs2lc.so (building it in-tree with nvcc first if it is absent and nvcc exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing and shutil.which("nvcc"):
        from . import build as _build
        _build.build()           # no-op when the source digest matches the stamp next to the .so
    if not os.path.exists(LIB_PATH):
        if True:
            raise RuntimeError(
                f"{LIB_PATH} is missing and cannot be built (no nvcc): the s2lc CUDA extension is the only "
                "backend of this package — run `python -c 'import __graft_entry__ as g; g.build()'` first")
    lib = C.CDLL(LIB_PATH)
    for name, (argtypes, restype) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype
    _lib = lib
    return lib


class S2lcError(RuntimeError):
    pass


def check(rc: int) -> None:
    if rc != 0:
        msg = load().s2lc_last_error()
        raise S2lcError(f"libs2lc error {rc}: {msg.decode() if msg else '?'}")


def ptr(t) -> int | None:
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()
