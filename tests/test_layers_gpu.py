"""Operator-level parity of the memory-bound forward layers (through the C ABI) against the CPU oracle.

maxpool is a selection, so the bf16 result must be bit-identical to the oracle's on bf16-rounded inputs
(maxpool_layer.cpp:258-288 == maxpool_layer_kernels.cu:58-101).  The stride-1 SPP windows take the
shared-memory slab kernel (separable row/column maxima); every other geometry the direct kernel.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle.pyoracle import O

pytestmark = pytest.mark.gpu


def _bf16_round(a):
    return torch.from_numpy(a).to(torch.bfloat16).float().numpy()


@pytest.mark.parametrize("size", [5, 9, 13, 3, 2])
@pytest.mark.parametrize("B,C,H,W", [(2, 32, 19, 19), (3, 24, 13, 13), (1, 8, 7, 11)])
def test_maxpool_stride1_slab_bit_exact(gpu, size, B, C, H, W):
    from darknet_b200 import ops
    rng = np.random.default_rng(size * 100 + C)
    x = _bf16_round(rng.standard_normal((B, C, H, W)).astype(np.float32))
    ref = O.maxpool(x, size, 1)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
    y = ops.maxpool(xd, C, size, 1)
    got = ops.to_nchw(y, C).cpu().numpy()
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)
    # the direct kernel (fp32 view) agrees with the slab kernel
    xf = ops.to_nhwc(torch.from_numpy(x).to(gpu), dtype=torch.float32)
    yf = ops.maxpool(xf, C, size, 1)
    assert np.array_equal(ops.to_nchw(yf, C).cpu().numpy(), ref)


def test_maxpool_slab_inside_concat_buffer(gpu):
    """input and output are channel slices of wider NHWC buffers (the SPP route's concat buffer)."""
    from darknet_b200 import ops, lib as L
    B, C, H, W, ldc = 2, 16, 19, 19, 64
    rng = np.random.default_rng(7)
    x = _bf16_round(rng.standard_normal((B, C, H, W)).astype(np.float32))
    buf = torch.full((B, H, W, ldc), 7.0, dtype=torch.bfloat16, device=gpu)
    buf[..., 48:64] = torch.from_numpy(x).to(gpu).permute(0, 2, 3, 1).to(torch.bfloat16)
    xv = L.view(buf, B, H, W, C, ldc, c_off=48)
    yv = L.view(buf, B, H, W, C, ldc, c_off=16)
    L.check(L.lib().yc_maxpool_fwd(xv, yv, 9, 1, 8, 0, L.stream_ptr()), "yc_maxpool_fwd")
    torch.cuda.synchronize()
    ref = O.maxpool(x, 9, 1)
    got = buf[..., 16:32].float().permute(0, 3, 1, 2).cpu().numpy()
    assert np.array_equal(got, ref)
    assert torch.all(buf[..., :16] == 7.0) and torch.all(buf[..., 32:48] == 7.0)


@pytest.mark.parametrize("size,stride", [(2, 2), (3, 2)])
def test_maxpool_strided_direct(gpu, size, stride):
    from darknet_b200 import ops
    rng = np.random.default_rng(3)
    x = _bf16_round(rng.standard_normal((2, 40, 26, 26)).astype(np.float32))
    ref = O.maxpool(x, size, stride)
    y = ops.maxpool(ops.to_nhwc(torch.from_numpy(x).to(gpu)), 40, size, stride)
    assert np.array_equal(ops.to_nchw(y, 40).cpu().numpy(), ref)


@pytest.mark.parametrize("gaussian", [False, True])
@pytest.mark.parametrize("H,W,classes,dtype", [(19, 19, 80, torch.float32), (13, 17, 3, torch.bfloat16), (76, 76, 80, torch.float32)])
def test_yolo_decode_op(gpu, gaussian, H, W, classes, dtype):
    """ForwardYoloLayerGpu's activation pass (yolo_layer.cpp:388-404 / gaussian_yolo_layer.cpp:497-516): 1e-6 absolute."""
    from darknet_b200 import ops
    rng = np.random.default_rng(11)
    n = 3
    E = (9 if gaussian else 5) + classes
    C = n * E
    x = (rng.standard_normal((2, C, H, W)) * 3).astype(np.float32)
    if dtype == torch.bfloat16:
        x = _bf16_round(x)
    ref = O.yolo_decode(x, n, classes, 1.2, gaussian=gaussian)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu), dtype=dtype)
    got = ops.yolo_decode(xd, C, n, classes, 1.2, gaussian=gaussian).cpu().numpy()
    assert np.abs(got - ref).max() <= 1e-6


@pytest.mark.parametrize("B,H,W,stride", [(2, 16, 20, 1), (3, 13, 12, 1), (1, 7, 44, 1), (2, 9, 10, 1), (2, 16, 16, 2), (1, 13, 11, 2)])
def test_stem_im2col_operand_bit_exact(gpu, B, H, W, stride):
    """yc_stem_im2col builds the first conv's GEMM operand from the fp32 NCHW image: row = output pixel, column
    (kh*3 + kw)*3 + ci, zero padding, columns 27..31 zero, bf16 round-to-nearest.  Stride 1 with W % 4 == 0 takes the
    four-pixels-per-thread kernel, everything else the per-pixel kernel; both must equal the numpy statement bit for bit
    (warp tails included: B*H*W is not a multiple of 32 / 128 in most cases)."""
    from darknet_b200 import lib as L
    rng = np.random.default_rng(B * 1000 + H * 10 + W)
    x = rng.standard_normal((B, 3, H, W)).astype(np.float32)
    OH, OW = (H + 2 - 3) // stride + 1, (W + 2 - 3) // stride + 1
    xp = np.pad(x, ((0, 0), (0, 0), (1, 1), (1, 1)))
    ref = np.zeros((B, OH, OW, 32), np.float32)
    for kh in range(3):
        for kw in range(3):
            for ci in range(3):
                ref[..., (kh * 3 + kw) * 3 + ci] = xp[:, ci, kh:kh + (OH - 1) * stride + 1:stride, kw:kw + (OW - 1) * stride + 1:stride]
    ref = _bf16_round(ref)
    xd = torch.from_numpy(x).to(gpu)
    out = torch.full((B, OH, OW, 32), 7.0, device=gpu, dtype=torch.bfloat16)
    t = L.view(out, B, OH, OW, 32)
    L.check(L.lib().yc_stem_im2col(xd.data_ptr(), 3, H, W, 3, stride, 1, C.byref(t), L.stream_ptr()), "yc_stem_im2col")
    torch.cuda.synchronize()
    assert np.array_equal(out.float().cpu().numpy(), ref)
