"""ctypes binding of the C-ABI kernel library (``include/unet_zoo_b200.h``).

The shared library is built in-tree by ``__graft_entry__.build()`` / ``tools/build.sh``.  There is no CPU path and no
fallback: if the library is missing or a call fails, a ``RuntimeError`` is raised.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# UZ_LIB_PATH: load another build of the same library (A/B timing of two builds inside one GPU session)
LIB_PATH = os.environ.get("UZ_LIB_PATH") or os.path.join(_HERE, "libunet_zoo_b200.so")

# mode / flag constants (mirrors include/unet_zoo_b200.h)
CONV3X3, CONV3X3_DGRAD, CONV1X1, CONV3X3_S2, CONV3X3_S2_DGRAD = 0, 1, 2, 3, 4
CONV1X1_S2, CONV1X1_S2_DGRAD, CONVT2X2, CONVT2X2_DGRAD, UP2X_CONV3X3 = 5, 6, 7, 8, 9
UP2X_CONV3X3_DGRAD = 10
FLAG_RELU, FLAG_ACCUMULATE, FLAG_STATS = 1, 2, 4
FLAG_RES1X1, FLAG_RES1X1_S2 = 8, 16


def dilation(d: int) -> int:
    """UZ_DILATION(d): OR into CONV3X3 / CONV3X3_DGRAD for Conv2d(3, padding = dilation = d); 0 for d = 1."""
    return (int(d) << 8) if d > 1 else 0

_P = c_void_p
_PROTOS = {
    "uz_last_error": (c_char_p, []),
    "uz_version": (c_int, []),
    "uz_device_sm_count": (c_int, []),
    "uz_conv_stats_rows": (c_int, []),
    "uz_reduce_rows": (c_int, []),
    "uz_up2x_phase_form": (c_int, [c_int, c_int, c_int]),
    "uz_pack_up2x": (c_int, [_P, _P, c_int, c_int, c_int, _P]),
    "uz_pack_weight": (c_int, [_P, _P, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong, c_longlong, _P, _P]),
    "uz_pack_weights_batched": (c_int, [_P, c_int, c_longlong, _P]),
    "uz_stem_im2col": (c_int, [_P, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "uz_nchw_to_nhwc": (c_int, [_P, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "uz_conv_igemm": (c_int, [c_int, c_int, c_int, c_int, c_int, _P, _P, _P, _P, c_int, _P, _P, c_int, c_int, c_int,
                              _P, c_int, _P]),
    "uz_conv_wgrad_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "uz_conv_wgrad": (c_int, [c_int, c_int, c_int, c_int, _P, c_int, c_int, _P, c_int, c_int, _P, c_int, c_int, c_int,
                              _P, c_size_t, _P]),
    "uz_bn_stats": (c_int, [_P, c_int, c_longlong, c_int, _P, _P]),
    "uz_bn_finalize": (c_int, [_P, c_int, c_int, c_longlong, _P, _P, c_float, c_float, _P, _P, _P, _P, _P]),
    "uz_bn_apply": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_int, _P, c_int, _P, c_int, _P, c_int, _P]),
    "uz_bn_bwd_workspace_bytes": (c_size_t, [c_int]),
    "uz_bn_bwd": (c_int, [_P, c_int, _P, c_int, _P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P, c_int, _P, _P, _P,
                          c_int, _P, c_size_t, _P]),
    "uz_colsum_workspace_bytes": (c_size_t, [c_int]),
    "uz_colsum": (c_int, [_P, c_int, c_longlong, c_int, _P, c_int, _P, c_size_t, _P]),
    "uz_colsum_static_max_channels": (c_int, []),
    "uz_colsum_static": (c_int, [_P, c_int, c_longlong, c_int, _P, c_int, _P]),
    "uz_gate_psi_fwd": (c_int, [_P, c_int, _P, c_int, c_longlong, c_int, _P, _P, _P, _P, _P, _P, _P]),
    "uz_gate_apply": (c_int, [_P, c_int, _P, _P, c_longlong, c_int, _P, c_int, _P]),
    "uz_gate_bwd_workspace_bytes": (c_size_t, [c_longlong, c_int]),
    "uz_gate_bwd": (c_int, [_P, c_int, _P, c_int, _P, _P, c_longlong, c_int, _P, c_int, c_int, _P, c_int, _P, c_int, c_int,
                            _P, _P, _P, _P, c_int, _P, _P, _P, _P, _P, c_size_t, _P]),
    "uz_bilinear2x_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "uz_addpe_fwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, _P, c_int, _P, c_int, _P]),
    "uz_addpe_bwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, _P, c_int, _P, c_int, _P, c_int, c_int, _P]),
    "uz_adaptive_avgpool_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int, c_int, _P]),
    "uz_adaptive_avgpool_bwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int, _P]),
    "uz_bilinear_resize_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int, c_int, c_int, _P]),
    "uz_bilinear_resize_bwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "uz_bilinear_resize_bwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int, c_int, _P,
                                       c_size_t, _P]),
    "uz_residual_add_fwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P, c_int, c_int, _P]),
    "uz_nhwc_to_nchw_f32": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P]),
    "uz_dual_bn_bwd_workspace_bytes": (c_size_t, [c_int]),
    "uz_dual_bn_add_relu_fwd": (c_int, [_P, c_int, _P, c_int, _P, c_int, _P, c_int, c_int, c_int, c_int, _P, c_int, _P, _P]),
    "uz_dual_bn_add_relu_bwd": (c_int, [_P, c_int, _P, c_int, _P, c_int, _P, c_int, c_int, c_int, c_int, _P, c_int, _P, _P,
                                        _P, c_int, c_int, _P, _P, c_int, c_int, _P, _P, c_size_t, _P]),
    "uz_bn_bwd_coef": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P, _P, _P, _P, c_int, _P,
                               c_size_t, _P]),
    "uz_side_head_bwd_workspace_bytes": (c_size_t, [c_int, c_int]),
    "uz_side_head_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, c_int, _P, _P, _P, _P]),
    "uz_side_head_bwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_int, _P, _P, _P, c_size_t, _P]),
    "uz_residual_add_bwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P, c_int, _P, c_int, c_int, _P, c_int,
                                    c_int, c_int, _P]),
    "uz_bmm_tn": (c_int, [_P, c_longlong, c_longlong, _P, c_longlong, _P, c_longlong, c_longlong, c_int, c_int, c_int,
                          c_int, _P]),
    "uz_softmax_rows": (c_int, [_P, c_longlong, _P, c_longlong, c_longlong, c_int, c_float, _P]),
    "uz_softmax_rows_bwd": (c_int, [_P, c_longlong, _P, c_longlong, _P, c_longlong, c_longlong, c_int, c_float, _P]),
    "uz_softmax_cols": (c_int, [_P, _P, c_int, c_int, c_int, c_float, _P]),
    "uz_softmax_cols_bwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_float, _P]),
    "uz_transpose_bf16": (c_int, [_P, c_longlong, c_longlong, _P, c_longlong, c_longlong, c_int, c_int, c_int, _P]),
    "uz_resize_bilinear_u8": (c_int, [_P, c_int, c_int, c_int, c_int, _P, _P, c_int, c_int, _P, _P, c_int, _P, _P, c_int,
                                      _P]),
    "uz_pad_copy": (c_int, [_P, c_longlong, c_int, c_int, _P, c_longlong, c_int, c_int, c_int, c_int, _P]),
    "uz_nearest2x_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "uz_bilinear2x_bwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int, _P]),
    "uz_head_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, c_int, _P, _P, _P, _P]),
    "uz_head_bwd_workspace_bytes": (c_size_t, [c_int]),
    "uz_head_bwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_int, _P, _P, c_int, _P, c_size_t,
                            _P]),
    "uz_bn_head_workspace_bytes": (c_size_t, [c_int, c_int]),
    "uz_bn_head_fwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_int, c_int, _P, _P, _P, _P]),
    "uz_bn_head_bwd": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, c_int, _P, _P, _P, c_int, _P, _P, _P,
                               _P, _P, c_int, _P, c_size_t, _P]),
    # fp32 parity mode (forward only, csrc/f32.cu)
    "uz_f32_pack_weight": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, c_longlong, c_longlong, _P, _P]),
    "uz_f32_conv": (c_int, [c_int, c_int, c_int, c_int, c_int, _P, _P, _P, _P, c_int, c_int, _P, _P, c_int, c_int, _P]),
    "uz_f32_nchw_to_nhwc": (c_int, [_P, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "uz_f32_nhwc_to_nchw": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P]),
    "uz_f32_bn_stats": (c_int, [_P, c_int, c_longlong, c_int, _P, _P]),
    "uz_f32_affine_act": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P, c_int, _P, c_int, _P, c_int, _P, c_int,
                                  _P]),
    "uz_f32_bilinear2x": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "uz_f32_gate_apply": (c_int, [_P, c_int, _P, _P, c_longlong, c_int, _P, c_int, _P]),
    # training-step tail (csrc/tail.cu)
    "uz_tail_workspace_bytes": (c_size_t, []),
    "uz_bce_dice": (c_int, [_P, _P, c_longlong, c_float, c_float, _P, _P, _P, c_size_t, _P]),
    "uz_adamw_flat": (c_int, [_P, _P, _P, _P, c_longlong, _P, _P, _P, c_size_t, _P]),
    # input pipeline tail (csrc/input.cu)
    "uz_stem_im2col_u8": (c_int, [_P, c_int, c_int, c_int, c_int, _P, _P, _P, c_int, _P]),
    "uz_mask_u8_to_f32": (c_int, [_P, c_longlong, _P, _P]),
}

_lib = None


def exported_symbols():
    """Names every build of the library must export (checked by the CPU test-suite against the header)."""
    return sorted(_PROTOS)


def load():
    """Load the kernel library, raising a RuntimeError that says how to build it when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"unet_zoo_b200: kernel library {LIB_PATH} not found - build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().uz_last_error()
        raise RuntimeError(f"unet_zoo_b200: {what} failed (rc={rc}): {msg.decode() if msg else '?'}")
