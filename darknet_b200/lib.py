"""ctypes binding of libyc_b200.so (the C ABI declared in include/yc_b200.h).

There is no fallback: if the shared library is missing or a call fails, this raises.
PyTorch is used by callers only for device memory and streams (tensor.data_ptr(), current stream).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libyc_b200.so")

LOGISTIC, RELU, LINEAR, LEAKY, SWISH, MISH = 0, 1, 4, 8, 16, 17
ACT = {"logistic": 0, "relu": 1, "linear": 4, "leaky": 8, "swish": 16, "mish": 17}
BF16, F32 = 0, 1


class YcError(RuntimeError):
    pass


class Tensor(C.Structure):
    """struct yc_tensor: NHWC view."""
    _fields_ = [("ptr", C.c_void_p), ("n", C.c_int), ("h", C.c_int), ("w", C.c_int), ("c", C.c_int),
                ("ldc", C.c_int), ("dtype", C.c_int)]


class ConvDesc(C.Structure):
    _fields_ = [("ksize", C.c_int), ("stride", C.c_int), ("pad", C.c_int), ("dilation", C.c_int),
                ("activation", C.c_int), ("residual_activation", C.c_int)]


class YoloLossDesc(C.Structure):
    _fields_ = [("anchors", C.c_int), ("total", C.c_int), ("classes", C.c_int), ("max_boxes", C.c_int),
                ("net_w", C.c_int), ("net_h", C.c_int), ("iou_loss", C.c_int), ("iou_thresh_kind", C.c_int),
                ("focal_loss", C.c_int), ("ignore_thresh", C.c_float), ("truth_thresh", C.c_float),
                ("iou_thresh", C.c_float), ("iou_normalizer", C.c_float), ("cls_normalizer", C.c_float),
                ("label_smooth_eps", C.c_float), ("max_delta", C.c_float), ("gaussian", C.c_int),
                ("uc_normalizer", C.c_float)]


class Detection(C.Structure):
    """struct yc_detection == the reference's struct Detection (yolo_core.h:598-609) on the path."""
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("w", C.c_float), ("h", C.c_float), ("classes", C.c_int),
                ("prob", C.POINTER(C.c_float)), ("objectness", C.c_float), ("sort_class", C.c_int),
                ("uc", C.c_float * 4)]


NET_FUSE_BN, NET_FUSE_SHORTCUT, NET_ELIDE_ROUTE, NET_CUDA_GRAPH, NET_PULL_HEADS, NET_REF_CPU_MAXPOOL = 1, 2, 4, 8, 16, 32
NET_STEM_IM2COL = 512
NET_NO_STEM_PAIR, NET_NO_PAIR_S2, NET_NO_PAIR_FUSE, NET_PLAN_SHIFT = 1024, 2048, 4096, 16
NET_STEM_DIRECT, NET_PDL = 16384, 32768
NET_STEM_OPERAND = -2147483648  # bit 31 of the int flags word
PLAN_NO_HALO, PLAN_FORCE_HALO, PLAN_NO_HALO_PAIR, PLAN_NO_W_RES, PLAN_W_RES_1X1, PLAN_ONE_OUT_BUF, PLAN_COST_V1 = 1, 2, 4, 8, 16, 32, 64
PLAN_TWO_OUT_BUFS, PLAN_GENERIC_EPI, PLAN_EPI_SETS_V1, PLAN_MISH_NEWTON, PLAN_NO_TILE_PAIR = 128, 256, 512, 1024, 2048
PLAN_NO_CTA_PAIR, PLAN_CTA_PAIR_ALL, PLAN_PDL = 4096, 8192, 16384
PLAN_HEAD_DECODE = 32  # (the bit of the former no-op PLAN_ONE_OUT_BUF)
# every plan switch off: the plainest kernel variant of each layer (parity coverage of the non-default paths)
NET_PLAIN_PLANS = NET_NO_STEM_PAIR | NET_NO_PAIR_S2 | NET_NO_PAIR_FUSE | ((PLAN_NO_HALO | PLAN_NO_W_RES | PLAN_GENERIC_EPI | PLAN_NO_TILE_PAIR | PLAN_NO_CTA_PAIR) << NET_PLAN_SHIFT)
NET_DEFAULT = 1 | 2 | 4 | 8 | 16 | 512
NET_TRAIN, NET_DGRAD_ACCUMULATE, NET_BN_BIAS_GRAD, NET_TRAIN_REF_CPU = 64, 128, 256, 8192

_lib = None

_P = C.c_void_p
_TP = C.POINTER(Tensor)
_SIGS = {
    "yc_last_error": (C.c_char_p, []),
    "yc_set_device": (C.c_int, [C.c_int]),
    "yc_get_device": (C.c_int, []),
    "yc_device_count": (C.c_int, []),
    "yc_sm_count": (C.c_int, []),
    "yc_malloc": (_P, [C.c_size_t]),
    "yc_free": (C.c_int, [_P]),
    "yc_malloc_host": (_P, [C.c_size_t]),
    "yc_free_host": (C.c_int, [_P]),
    "yc_push": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "yc_pull": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "yc_memset": (C.c_int, [_P, C.c_int, C.c_size_t, _P]),
    "yc_stream_sync": (C.c_int, [_P]),
    "yc_stream_create": (C.c_int, [C.POINTER(_P)]),
    "yc_stream_destroy": (C.c_int, [_P]),
    "yc_launch_count": (C.c_longlong, [C.c_int]),
    "yc_nchw_to_nhwc": (C.c_int, [_P, _TP, _P]),
    "yc_nhwc_to_nchw": (C.c_int, [_TP, _P, _P]),
    "yc_stem_im2col": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _TP, _P]),
    "yc_stem_conv_direct": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, _TP, _P]),
    "yc_stem_conv_tc": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, _TP, _P]),
    "yc_conv_padded_in_c": (C.c_int, [C.c_int]),
    "yc_conv_packed_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "yc_conv_pack": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "yc_conv_plan_create": (C.c_int, [C.POINTER(_P), C.POINTER(ConvDesc), _TP, _P, _P, _P, _TP, _TP]),
    "yc_conv_plan_create_ex": (C.c_int, [C.POINTER(_P), C.POINTER(ConvDesc), _TP, _P, _P, _P, _TP, _TP, _TP, C.c_uint]),
    "yc_conv_plan_destroy": (C.c_int, [_P]),
    "yc_conv_fprop": (C.c_int, [_P, _P]),
    "yc_conv_fprop_set_issue_v2": (C.c_int, [C.c_int]),
    "yc_conv_plan_info": (C.c_int, [_P, C.c_int]),
    "yc_conv_plan_set_stats": (C.c_int, [_P, _P, _P]),
    "yc_conv_plan_set_bias": (C.c_int, [_P, _P, _P]),
    "yc_conv_plan_set_yolo_decode": (C.c_int, [_P, _P, C.c_int, C.c_float, C.c_int]),
    "yc_conv_plan_set_bn_finalize": (C.c_int, [_P, C.c_longlong, _P, _P, _P, _P]),
    "yc_bn_finalize": (C.c_int, [_P, _P, C.c_longlong, _P, _P, _P, _P, C.c_int, _P]),
    "yc_bn_act_fwd": (C.c_int, [_TP, _P, _P, _P, _P, C.c_int, _TP, _P]),
    "yc_bn_bwd": (C.c_int, [_TP, _TP, _P, _P, _P, _P, C.c_int, _P, _P, _P, _TP, _P]),
    "yc_bias_act_bwd": (C.c_int, [_TP, _TP, C.c_int, _P, _P, _TP, _P]),
    "yc_bn_bwd_workspace_floats": (C.c_size_t, [C.c_int]),
    "yc_bias_act_bwd_noalloc": (C.c_int, [_TP, _TP, C.c_int, _P, _TP, _P]),
    "yc_bias_act_bwd_noalloc_max_channels": (C.c_size_t, []),
    "yc_conv_dgrad_channels": (C.c_int, [C.c_int]),
    "yc_conv_dgrad_packed_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "yc_conv_pack_dgrad": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "yc_conv_dgrad_dilated_size": (C.c_int, [C.POINTER(ConvDesc), C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "yc_conv_dgrad_dilate": (C.c_int, [_TP, C.c_int, _TP, _P]),
    "yc_conv_dgrad_plan_create": (C.c_int, [C.POINTER(_P), C.POINTER(ConvDesc), _TP, _P, _TP, C.c_int]),
    "yc_conv_wgrad_ld": (C.c_int, [C.c_int]),
    "yc_conv_wgrad_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "yc_conv_wgrad_plan_create": (C.c_int, [C.POINTER(_P), C.POINTER(ConvDesc), _TP, _TP, C.c_int, _P]),
    "yc_conv_wgrad_plan_destroy": (C.c_int, [_P]),
    "yc_conv_wgrad": (C.c_int, [_P, _P]),
    "yc_conv_wgrad_plan_info": (C.c_int, [_P, C.c_int]),
    "yc_conv_wgrad_set_tap_groups": (C.c_int, [C.c_int]),
    "yc_conv_wgrad_unpack": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "yc_conv_wgrad_pack": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "yc_maxpool_fwd_train": (C.c_int, [_TP, _TP, _P, C.c_int, C.c_int, C.c_int, _P]),
    "yc_maxpool_bwd": (C.c_int, [_TP, _P, _TP, C.c_int, C.c_int, C.c_int, _P]),
    "yc_upsample_bwd": (C.c_int, [_TP, _TP, C.c_int, C.c_float, _P]),
    "yc_axpy_view": (C.c_int, [_TP, _TP, _P]),
    "yc_shortcut_bwd": (C.c_int, [_TP, _TP, C.c_int, _TP, _TP, _P]),
    "yc_sgd_update_conv": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, C.c_float, _P, _P, _P, _P, _P, _P, _P]),
    "yc_maxpool_fwd": (C.c_int, [_TP, _TP, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "yc_maxpool_s1_multi": (C.c_int, [_TP, C.c_int, _P, _P, _P, _P]),
    "yc_maxpool_s1_multi_eligible": (C.c_int, [_TP, _TP, C.c_int, C.c_int, C.c_int]),
    "yc_upsample_fwd": (C.c_int, [_TP, _TP, C.c_int, C.c_float, _P]),
    "yc_shortcut_fwd": (C.c_int, [_TP, _TP, _TP, C.c_int, _P]),
    "yc_route_copy": (C.c_int, [_TP, _TP, _P]),
    "yc_scale_channels_fwd": (C.c_int, [_TP, _TP, _TP, C.c_int, _P]),
    "yc_reorg_fwd": (C.c_int, [_TP, _TP, C.c_int, C.c_int, _P]),
    "yc_scale_channels_bwd": (C.c_int, [_TP, _TP, C.c_int, _TP, _TP, _TP, _TP, _P]),
    "yc_reorg_bwd": (C.c_int, [_TP, _TP, C.c_int, C.c_int, _P]),
    "yc_yolo_decode": (C.c_int, [_TP, _P, C.c_int, C.c_int, C.c_float, C.c_int, _P]),
    "yc_yolo_loss_stats_floats": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "yc_yolo_loss": (C.c_int, [C.POINTER(YoloLossDesc), _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _TP, _P, _P]),
    "yc_yolo_extract": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, C.c_int,
                                  C.c_float, C.c_int, _P, C.c_int, _P, _P]),
    "yc_network_load": (C.c_int, [C.POINTER(_P), C.c_char_p, C.c_char_p, C.c_int, C.c_int]),
    "yc_network_load_text": (C.c_int, [C.POINTER(_P), C.c_char_p, C.c_char_p, C.c_int, C.c_int]),
    "yc_network_free": (C.c_int, [_P]),
    "yc_network_set_conv_weights": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P]),
    "yc_network_finalize": (C.c_int, [_P]),
    "yc_network_info": (C.c_int, [_P, C.c_int]),
    "yc_network_layer_info": (C.c_int, [_P, C.c_int, C.c_int]),
    "yc_network_layer_input": (C.c_int, [_P, C.c_int, C.c_int]),
    "yc_network_layer_float": (C.c_float, [_P, C.c_int, C.c_int]),
    "yc_network_layer_array": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_int]),
    "yc_network_flops": (C.c_double, [_P]),
    "yc_network_load_weights": (C.c_int, [_P, C.c_char_p]),
    "yc_network_save_weights": (C.c_int, [_P, C.c_char_p, C.c_int]),
    "yc_network_init_weights": (C.c_int, [_P, C.c_uint, C.c_int]),
    "yc_network_current_rate": (C.c_float, [_P, C.c_longlong, C.c_int]),
    "yc_network_set_schedule": (C.c_int, [_P, C.c_int, C.c_longlong]),
    "yc_network_resize": (C.c_int, [_P, C.c_int, C.c_int]),
    "yc_network_predict": (_P, [_P, _P]),
    "yc_network_predict_u8": (_P, [_P, _P, C.c_int]),
    "yc_network_predict_u8_resized": (_P, [_P, _P, C.c_int, C.c_int, C.c_int]),
    "yc_network_prefetch": (C.c_int, [_P, _P]),
    "yc_network_input_buffer_slot": (_P, [_P, C.c_int]),
    "yc_network_input_buffer_u8": (_P, [_P]),
    "yc_u8hwc_to_nchw": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "yc_resize_u8": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.c_int, C.c_int, _P]),
    "yc_resize_u8_to_nchw": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.c_int, C.c_int, _P]),
    "yc_network_input_buffer": (_P, [_P]),
    "yc_network_forward": (C.c_int, [_P, _P, _P]),
    "yc_network_layer_output": (C.c_int, [_P, C.c_int, _P]),
    "yc_network_yolo_output": (_P, [_P, C.c_int]),
    "yc_network_profile": (C.c_int, [_P, _P, C.c_int, _P]),
    "yc_network_time_ops": (C.c_int, [_P, _P, C.c_int, C.c_int, _P, _P]),
    "yc_network_train_datum": (C.c_int, [_P, _P, _P, C.POINTER(C.c_float)]),
    "yc_network_forward_backward": (C.c_int, [_P, _P, _P, _P]),
    "yc_network_update": (C.c_int, [_P, C.c_float, C.c_float, C.c_float, _P]),
    "yc_network_train_profile": (C.c_int, [_P, _P, _P, C.c_int, _P]),
    "yc_network_train_info": (C.c_int, [_P, C.c_int]),
    "yc_network_layer_delta": (C.c_int, [_P, C.c_int, _P]),
    "yc_network_layer_x": (C.c_int, [_P, C.c_int, _P]),
    "yc_network_get_param": (C.c_int, [_P, C.c_int, C.c_char_p, _P, C.c_int]),
    "yc_dp_unique_id": (C.c_int, [_P]),
    "yc_dp_nccl_version": (C.c_int, []),
    "yc_network_dp_init": (C.c_int, [_P, _P, C.c_int, C.c_int]),
    "yc_network_dp_info": (C.c_int, [_P, C.c_int]),
    "yc_network_dp_set_overlap": (C.c_int, [_P, C.c_int]),
    "yc_network_dp_reduce": (C.c_int, [_P, _P]),
    "yc_network_dp_discard_step": (C.c_int, [_P]),
    "yc_network_get_boxes": (C.POINTER(Detection), [_P, C.c_int, C.c_float, C.POINTER(C.c_int)]),
    "yc_free_detections": (None, [C.POINTER(Detection), C.c_int]),
    "yc_nms_sort": (None, [C.POINTER(Detection), C.c_int, C.c_int, C.c_float, C.c_int, C.c_float]),
    "yc_network_detect": (C.c_int, [_P, C.c_float, C.c_float, C.POINTER(C.POINTER(C.c_float)), _P, C.POINTER(C.c_int)]),
}


def declared_symbols():
    """Every function the public header declares (parsed from include/yc_b200.h)."""
    import re
    hdr = open(os.path.join(HERE, "..", "include", "yc_b200.h")).read()
    return sorted(set(re.findall(r"YC_API[^;(]*?\b(yc_\w+)\s*\(", hdr)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            raise YcError(f"{SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
        L = C.CDLL(SO)
        for name, (res, args) in _SIGS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise YcError(f"{what} failed ({rc}): {lib().yc_last_error().decode()}")


def view(t, n, h, w, c, ldc=None, c_off=0):
    """yc_tensor over a torch CUDA tensor's storage (NHWC). `t` is the buffer that holds the data."""
    import torch
    dtype = BF16 if t.dtype == torch.bfloat16 else F32
    assert t.dtype in (torch.bfloat16, torch.float32)
    ldc = c if ldc is None else ldc
    return Tensor(t.data_ptr() + c_off * t.element_size(), n, h, w, c, ldc, dtype)


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
