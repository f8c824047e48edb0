"""ctypes binding of ``libsoda_b200.so`` (the C ABI declared in ``include/soda_b200.h``).

This is the only door between the Python host side and the CUDA kernels.  There is no CPU
implementation behind any entry point: if the shared library is missing the import of this
module raises, and every call checks the returned status and raises ``SodaError`` with the
library's own message.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, os.environ.get("SODA_LIB_NAME", "libsoda_b200.so"))   # (A/B builds of the same ABI)

SODA_CONV_FP32 = 0
SODA_CONV_TC = 1
SODA_CONV_FLAG_INPUT_EXACT_TF32 = 1
SODA_CONV_FLAG_INPUT_EXACT_BF16 = 2
SODA_CONV_FLAG_OUT_HALF = 4
SODA_CONV_FLAG_OUT_PAIR = 8
SODA_ACT_F32 = 0
SODA_ACT_PAIR = 1
SODA_ACT_HALF = 2
SODA_PCONV_FLAG_INPUT_EXACT = 1


class SodaError(RuntimeError):
    """A C-ABI call returned a non-zero status."""


class ConvDesc(C.Structure):
    _fields_ = [
        ("N", C.c_int),
        ("H", C.c_int),
        ("W", C.c_int),
        ("Cin", C.c_int),
        ("Cout", C.c_int),
        ("k", C.c_int),
        ("stride", C.c_int),
        ("in_nchw", C.c_int),
        ("kind", C.c_int),
        ("flags", C.c_int),
        ("in_batch_stride", C.c_int64),
        ("in_pix_stride", C.c_int64),
        ("out_batch_stride", C.c_int64),
        ("out_pix_stride", C.c_int64),
        ("res_batch_stride", C.c_int64),
        ("res_pix_stride", C.c_int64),
    ]


class PConvDesc(C.Structure):
    _fields_ = [
        ("N", C.c_int),
        ("H", C.c_int),
        ("W", C.c_int),
        ("Cin", C.c_int),
        ("Cout", C.c_int),
        ("k", C.c_int),
        ("stride", C.c_int),
        ("in_fmt", C.c_int),
        ("out_fmt", C.c_int),
        ("res_fmt", C.c_int),
        ("flags", C.c_int),
        ("in_batch_stride", C.c_int64),
        ("in_pix_stride", C.c_int64),
        ("out_batch_stride", C.c_int64),
        ("out_pix_stride", C.c_int64),
        ("res_batch_stride", C.c_int64),
        ("res_pix_stride", C.c_int64),
    ]


class NeuronParams(C.Structure):
    _fields_ = [
        ("dt_tau_mem", C.c_float),
        ("dt_tau_syn", C.c_float),
        ("v_leak", C.c_float),
        ("v_th", C.c_float),
        ("v_reset", C.c_float),
        ("alpha", C.c_float),
        ("jump_first", C.c_int),
        ("detach_reset", C.c_int),
    ]


_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> (restype, argtypes); must list every symbol include/soda_b200.h declares
SIGNATURES = {
    "soda_abi_version": (_i, []),
    "soda_last_error": (C.c_char_p, []),
    "soda_launch_count": (_i64, []),
    "soda_bin_events": (_i, [_vp, _vp, _vp, _i, _i64, _i, _i, _i, _i, _i, _vp, _i64, _i64, _vp, _vp]),
    "soda_conv_packed_elems": (_i64, [_i, _i, _i, _i]),
    "soda_conv_pack_weights": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "soda_conv_tc_supported": (_i, [C.POINTER(ConvDesc), _vp]),
    "soda_conv2d_fwd": (_i, [C.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _vp, _vp]),
    "soda_conv_lif_fwd": (_i, [C.POINTER(ConvDesc), _vp, _vp, _vp, _vp, _i, C.POINTER(NeuronParams), _vp, _vp, _vp,
                                _vp, _vp, _vp]),
    "soda_pconv_packed_bytes": (_i64, [_i, _i, _i]),
    "soda_pconv_pack_weights": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "soda_pconv_supported": (_i, [C.POINTER(PConvDesc), _vp, _vp, _vp]),
    "soda_pconv_fwd": (_i, [C.POINTER(PConvDesc), _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "soda_pconv_lif_fwd": (_i, [C.POINTER(PConvDesc), _vp, _vp, _vp, _vp, _vp, _i, C.POINTER(NeuronParams), _vp, _vp,
                                 _vp, _vp, _vp, _vp, _vp]),
    "soda_conv_lif_pool_supported": (_i, [_i, _i, _i, _i, _i, _i, _i, _i]),
    "soda_conv_lif_pool_fwd": (_i, [_vp, _i, _i, _i64, _i64, _vp, _vp, _i, _i, _i, _i, _i, _i, C.POINTER(NeuronParams), _vp,
                                    _vp, _vp, _i, _i64, _i64, _vp, _vp, _vp]),
    "soda_im2col3_frames": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i64, _i64, _vp]),
    "soda_act_convert": (_i, [_vp, _i, _i64, _i64, _vp, _i, _i64, _i64, _i64, _i64, _i, _vp, _vp]),
    "soda_lif_fmt_fwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i64, _i, _vp, _i, _i64, _vp, _i, _i64,
                              _vp, _vp, _vp, _vp]),
    "soda_lif_fwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "soda_lif_res_fwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i64, _i, _vp, _i64, _vp, _i64, _vp, _vp,
                              _vp]),
    "soda_lif_pool_fwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _i64, _vp, _vp,
                               _vp]),
    "soda_li_fwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp]),
    "soda_lif_bwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _vp, _i, _i64, _vp, _vp, _vp, _vp]),
    "soda_lif_bwd_strided": (_i, [C.POINTER(NeuronParams), _vp, _i, _i64, _vp, _vp, _vp, _i, _i64, _vp, _vp, _vp, _vp]),
    "soda_li_bwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i64, _vp, _vp, _vp, _vp]),
    "soda_pool2d_fwd": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i64, _i64, _vp, _i64, _i64, _vp]),
    "soda_channel_affine": (_i, [_vp, _i64, _i, _i64, _vp, _vp, _vp, _i64, _vp]),
    "soda_add_n": (_i, [C.POINTER(_vp), _i, _i64, _vp, _vp]),
    "soda_multibox_workspace_bytes": (_i64, [_i, _i]),
    "soda_multibox_detection": (_i, [_vp, _vp, _vp, _i, _i, _i, _f, _f, _vp, _vp, _vp]),
    "soda_multibox_static_max_bytes": (_i64, []),
    "soda_multibox_detection_static": (_i, [_vp, _vp, _vp, _i, _i, _i, _f, _f, _vp, _vp]),
    "soda_sli_fwd": (_i, [C.POINTER(NeuronParams), _vp, _vp, _vp, _i, _i64, _vp, _vp, _vp, _vp, _vp]),
    "soda_synapse_fwd": (_i, [_f, _f, _f, _f, _vp, _vp, _i, _i64, _vp, _vp, _vp]),
    "soda_lstm_gates": (_i, [_vp, _vp, _i64, _i, _vp, _vp, _vp]),
    "soda_unary_fwd": (_i, [_vp, _i64, _i, _vp, _vp]),
    "soda_upsample_nearest": (_i, [_vp, _i64, _i, _i, _i, _i, _vp, _vp]),
    "soda_roi_assign": (_i, [_vp, _vp, _i, _i, _i, _f, _vp, _vp, _vp, _vp]),
    "soda_bn_train_workspace_bytes": (_i64, [_i, _i]),
    "soda_bn_train_fwd": (_i, [_vp, _i, _i64, _i, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "soda_bn_train_bwd": (_i, [_vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "soda_bn_train_stats": (_i, [_vp, _vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp]),
    "soda_bn_train_fwd_from_sums": (_i, [_vp, _vp, _i64, _i, _i64, _i, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "soda_bn_train_bwd_from_sums": (_i, [_vp, _vp, _vp, _i64, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
}

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C snn_for_object_detection_b200/csrc`).  There is no CPU fallback."
    )

lib = C.CDLL(LIB_PATH)
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header/library mismatch
    _fn.restype = _res
    _fn.argtypes = _args


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib.soda_last_error()
        raise SodaError(f"{what} failed (status {status}): {msg.decode() if msg else '?'}")


def launch_count() -> int:
    return int(lib.soda_launch_count())
