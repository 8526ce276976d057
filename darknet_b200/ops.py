"""Per-operator host wrappers over the C ABI, taking torch CUDA tensors in NHWC.

These mirror the reference's per-layer GPU entry points (ForwardConvolutionalLayerGpu,
ForwardMaxpoolLayerGpu, ... — see include/yc_b200.h for file:line) one call at a time, for unit
tests and microbenchmarks.  The whole-network path lives in csrc/network.cu (NetworkPredict).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as L


def _chan_pad(c, dtype):
    e = 4 if dtype == torch.float32 else 8
    return (c + e - 1) // e * e


def to_nhwc(x_nchw: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """fp32 NCHW (device) -> NHWC [n,h,w,ldc] with the pixel pitch padded to 16 bytes."""
    n, c, h, w = x_nchw.shape
    ldc = _chan_pad(c, dtype)
    out = torch.zeros((n, h, w, ldc), dtype=dtype, device=x_nchw.device)
    x = x_nchw.contiguous().float()
    L.check(L.lib().yc_nchw_to_nhwc(x.data_ptr(), L.view(out, n, h, w, c, ldc), L.stream_ptr()), "yc_nchw_to_nhwc")
    return out


def to_nchw(x_nhwc: torch.Tensor, c: int) -> torch.Tensor:
    n, h, w, ldc = x_nhwc.shape
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x_nhwc.device)
    L.check(L.lib().yc_nhwc_to_nchw(L.view(x_nhwc, n, h, w, c, ldc), out.data_ptr(), L.stream_ptr()), "yc_nhwc_to_nchw")
    return out


class Conv:
    """One convolution layer: packed weights + plan over fixed buffers."""

    def __init__(self, x_buf, x_shape, w_oihw, bias, k, stride, pad, act, dil=1, scale=None,
                 out_dtype=torch.bfloat16, residual=None, residual_act="linear", x_ldc=None, x_coff=0,
                 y_buf=None, y_ldc=None, y_coff=0):
        lib = L.lib()
        n, h, w, c = x_shape
        out_c = w_oihw.shape[0]
        assert w_oihw.shape[1] == c and w_oihw.shape[2] == k
        dev = x_buf.device
        self.w32 = w_oihw.contiguous().float().to(dev)
        self.packed = torch.zeros(lib.yc_conv_packed_bytes(out_c, c, k) // 2, dtype=torch.bfloat16, device=dev)
        L.check(lib.yc_conv_pack(self.w32.data_ptr(), out_c, c, k, self.packed.data_ptr(), L.stream_ptr()), "yc_conv_pack")
        self.bias = bias.contiguous().float().to(dev)
        self.scale = None if scale is None else scale.contiguous().float().to(dev)
        eff = dil * (k - 1) + 1
        ho, wo = (h + 2 * pad - eff) // stride + 1, (w + 2 * pad - eff) // stride + 1
        if y_buf is None:
            y_ldc = _chan_pad(out_c, out_dtype)
            y_buf = torch.zeros((n, ho, wo, y_ldc), dtype=out_dtype, device=dev)
        self.y = y_buf
        self.out_shape = (n, ho, wo, out_c)
        x_ldc = x_ldc if x_ldc is not None else x_buf.shape[-1]
        xv = L.view(x_buf, n, h, w, c, x_ldc, x_coff)
        yv = L.view(y_buf, n, ho, wo, out_c, y_ldc, y_coff)
        rv = None
        if residual is not None:
            self.residual = residual
            rv = C.byref(L.view(residual, n, ho, wo, out_c, residual.shape[-1]))
        desc = L.ConvDesc(k, stride, pad, dil, L.ACT[act] if isinstance(act, str) else act,
                          L.ACT[residual_act] if isinstance(residual_act, str) else residual_act)
        self.plan = C.c_void_p()
        L.check(lib.yc_conv_plan_create(C.byref(self.plan), C.byref(desc), C.byref(xv), self.packed.data_ptr(),
                                        self.bias.data_ptr(), None if self.scale is None else self.scale.data_ptr(),
                                        rv, C.byref(yv)), "yc_conv_plan_create")

    def info(self):
        lib = L.lib()
        keys = ["block_n", "block_k", "stages", "tiles", "grid", "a_mode", "smem"]
        return {k: lib.yc_conv_plan_info(self.plan, i) for i, k in enumerate(keys)}

    def __call__(self):
        L.check(L.lib().yc_conv_fprop(self.plan, L.stream_ptr()), "yc_conv_fprop")
        return self.y

    def __del__(self):
        try:
            if getattr(self, "plan", None):
                L.lib().yc_conv_plan_destroy(self.plan)
                self.plan = None
        except Exception:
            pass


def maxpool(x, c, size, stride, pad=None, quirk=False):
    n, h, w, ldc = x.shape
    pad = size - 1 if pad is None else pad
    ho, wo = (h + pad - size) // stride + 1, (w + pad - size) // stride + 1
    y = torch.zeros((n, ho, wo, ldc), dtype=x.dtype, device=x.device)
    L.check(L.lib().yc_maxpool_fwd(L.view(x, n, h, w, c, ldc), L.view(y, n, ho, wo, c, ldc), size, stride, pad,
                                   int(quirk), L.stream_ptr()), "yc_maxpool_fwd")
    return y


def upsample(x, c, stride=2, scale=1.0):
    n, h, w, ldc = x.shape
    y = torch.zeros((n, h * stride, w * stride, ldc), dtype=x.dtype, device=x.device)
    L.check(L.lib().yc_upsample_fwd(L.view(x, n, h, w, c, ldc), L.view(y, n, h * stride, w * stride, c, ldc), stride,
                                    scale, L.stream_ptr()), "yc_upsample_fwd")
    return y


def shortcut(a, b, c, act="linear"):
    n, h, w, ldc = a.shape
    y = torch.zeros_like(a)
    L.check(L.lib().yc_shortcut_fwd(L.view(a, n, h, w, c, ldc), L.view(b, n, h, w, c, b.shape[-1]),
                                    L.view(y, n, h, w, c, ldc), L.ACT[act], L.stream_ptr()), "yc_shortcut_fwd")
    return y


def route_copy(src, c, dst, dst_coff, src_coff=0):
    n, h, w, ldc = src.shape
    L.check(L.lib().yc_route_copy(L.view(src, n, h, w, c, ldc, src_coff), L.view(dst, n, h, w, c, dst.shape[-1], dst_coff),
                                  L.stream_ptr()), "yc_route_copy")
    return dst


def scale_channels(x, s, c, act="linear"):
    n, h, w, ldc = x.shape
    y = torch.zeros_like(x)
    L.check(L.lib().yc_scale_channels_fwd(L.view(x, n, h, w, c, ldc), L.view(s, n, 1, 1, c, s.shape[-1]),
                                          L.view(y, n, h, w, c, ldc), L.ACT[act], L.stream_ptr()), "yc_scale_channels_fwd")
    return y


def reorg(x, c, stride, forward=0):
    n, h, w, ldc = x.shape
    if forward:
        oc, oh, ow = c // (stride * stride), h * stride, w * stride
    else:
        oc, oh, ow = c * stride * stride, h // stride, w // stride
    oldc = _chan_pad(oc, x.dtype)
    y = torch.zeros((n, oh, ow, oldc), dtype=x.dtype, device=x.device)
    L.check(L.lib().yc_reorg_fwd(L.view(x, n, h, w, c, ldc), L.view(y, n, oh, ow, oc, oldc), stride, forward,
                                 L.stream_ptr()), "yc_reorg_fwd")
    return y, oc


def yolo_decode(x, c, anchors, classes, scale_x_y, gaussian=False):
    n, h, w, ldc = x.shape
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    L.check(L.lib().yc_yolo_decode(L.view(x, n, h, w, c, ldc), out.data_ptr(), anchors, classes, scale_x_y,
                                   int(gaussian), L.stream_ptr()), "yc_yolo_decode")
    return out


def yolo_extract(decoded, anchors_wh, mask, classes, net_w, net_h, thresh, rows_per_image, gaussian=False):
    n, c, h, w = decoded.shape
    dev = decoded.device
    a = torch.as_tensor(anchors_wh, dtype=torch.float32, device=dev).contiguous()
    m = torch.as_tensor(mask, dtype=torch.int32, device=dev).contiguous()
    rows = torch.zeros((n, rows_per_image, 6 + classes), dtype=torch.float32, device=dev)
    counts = torch.zeros((n,), dtype=torch.int32, device=dev)
    L.check(L.lib().yc_yolo_extract(decoded.data_ptr(), n, h, w, len(mask), classes, a.data_ptr(), m.data_ptr(), net_w,
                                    net_h, thresh, int(gaussian), rows.data_ptr(), rows_per_image, counts.data_ptr(),
                                    L.stream_ptr()), "yc_yolo_extract")
    return rows, counts
