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


@pytest.mark.parametrize("gaussian,classes,Cin,B,H,W", [(False, 80, 256, 3, 19, 19), (False, 3, 64, 2, 13, 17), (True, 80, 128, 2, 38, 38),
                                                        (False, 80, 512, 5, 76, 76)])
def test_head_conv_with_fused_yolo_decode(gpu, gaussian, classes, Cin, B, H, W):
    """yc_conv_plan_set_yolo_decode (ForwardYoloLayerGpu inside the head convolution's epilogue) == the head convolution
    followed by yc_yolo_decode to 1e-6 absolute (the same accumulators; pass-through entries bit for bit); sizes cover several
    row tiles per image, tiles that straddle two images and the channel tail of the last 32-column group."""
    from darknet_b200 import lib as L, ops
    rng = np.random.default_rng(17)
    n = 3
    E = (9 if gaussian else 5) + classes
    N = n * E
    x = _bf16_round(rng.standard_normal((B, Cin, H, W)).astype(np.float32))
    w = _bf16_round((rng.standard_normal((N, Cin, 1, 1)) * (2.0 / Cin) ** 0.5 * 3).astype(np.float32))
    b = rng.uniform(-1, 1, N).astype(np.float32)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
    conv = ops.Conv(xd, (B, H, W, Cin), torch.from_numpy(w), torch.from_numpy(b), 1, 1, 0, "linear", out_dtype=torch.float32)
    want = ops.yolo_decode(conv(), N, n, classes, 1.05, gaussian=gaussian)
    fused = ops.Conv(xd, (B, H, W, Cin), torch.from_numpy(w), torch.from_numpy(b), 1, 1, 0, "linear", out_dtype=torch.float32)
    got = torch.full((B, N, H, W), -7.0, dtype=torch.float32, device=gpu)
    L.check(L.lib().yc_conv_plan_set_yolo_decode(fused.plan, got.data_ptr(), E, 1.05, int(gaussian)), "yc_conv_plan_set_yolo_decode")
    fused()
    torch.cuda.synchronize()
    # same accumulators; the logistic is the MUFU form (ex2 + rcp) here and libm's expf + a division in the separate kernel
    assert float((got - want).abs().max()) <= 1e-6, f"max diff {float((got - want).abs().max()):.3e}"
    e_w, e_h = (4, 6) if gaussian else (2, 3)   # the entries no logistic touches (gaussian_yolo_layer.cpp:497-516: mu_w, mu_h)
    assert torch.equal(got[:, e_w::E], want[:, e_w::E]) and torch.equal(got[:, e_h::E], want[:, e_h::E]), "w / h entries are passed through"


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


@pytest.mark.parametrize("act", ["linear", "leaky", "logistic"])
@pytest.mark.parametrize("B,C,H,W", [(2, 64, 13, 13), (3, 24, 5, 7), (1, 256, 19, 19)])
def test_scale_channels_fwd_bwd(gpu, act, B, C, H, W):
    """ForwardScaleChannelsLayerGpu / BackwardScaleChannelsLayerGpu (scale_channels_layer.c:134-160) vs the restatement
    (pinned to the reference by tests/test_oracle_pin.py).  Inputs are bf16-representable, so what is left is the bf16
    rounding of the stored results: 1e-2 relative; the channel sums are fp32 in a fixed order."""
    from darknet_b200 import ops
    rng = np.random.default_rng(C + H)
    r16 = lambda a: torch.from_numpy(a).to(torch.bfloat16).float().numpy()
    x = r16(rng.standard_normal((B, C, H, W)).astype(np.float32))
    s = r16(rng.uniform(0.1, 2.0, (B, C)).astype(np.float32))
    dy = r16(rng.standard_normal((B, C, H, W)).astype(np.float32))
    ds0 = r16(rng.standard_normal((B, C)).astype(np.float32))
    dx0 = r16(rng.standard_normal((B, C, H, W)).astype(np.float32))
    y_ref = O.scale_channels(x, s, act)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
    sd = ops.to_nhwc(torch.from_numpy(s.reshape(B, C, 1, 1)).to(gpu))
    yd = ops.scale_channels(xd, sd, C, act)
    y = ops.to_nchw(yd, C).cpu().numpy()
    assert np.abs(y - y_ref).max() <= 1e-2 * np.abs(y_ref).max()
    # backward on the GPU's own (bf16) output, like the layer does
    d_ref, ds_ref, dx_ref = O.scale_channels_backward(dy, y, x, s, act, ds0, dx0)
    dyd = ops.to_nhwc(torch.from_numpy(dy).to(gpu))
    dsd = ops.to_nhwc(torch.from_numpy(ds0.reshape(B, C, 1, 1)).to(gpu))
    dxd = ops.to_nhwc(torch.from_numpy(dx0).to(gpu))
    ops.scale_channels_bwd(dyd, yd, xd, sd, dsd, dxd, C, act)
    torch.cuda.synchronize()
    d = ops.to_nchw(dyd, C).cpu().numpy()
    ds = ops.to_nchw(dsd, C).cpu().numpy().reshape(B, C)
    dx = ops.to_nchw(dxd, C).cpu().numpy()
    assert np.abs(d - d_ref).max() <= 1e-2 * np.abs(d_ref).max()
    assert np.abs(dx - dx_ref).max() <= 1e-2 * np.abs(dx_ref).max()
    # ds: the kernel sums the bf16-rounded dy*act' in fp32 and stores bf16
    assert np.abs(ds - ds_ref).max() <= 1.5e-2 * np.abs(ds_ref).max()


@pytest.mark.parametrize("stride", [2, 4])
@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_reorg_fwd_bwd_bit_exact(gpu, stride, dtype):
    """ForwardReorgLayerGpu / BackwardReorgLayerGpu (reorg_layer.c:113-143): pure data movement, bit-exact against
    reorg_cpu's index math (fp32 tensors: exact; bf16 tensors: exact on bf16-representable inputs)."""
    from darknet_b200 import ops
    rng = np.random.default_rng(stride)
    B, C, H, W = 2, 6, 8 * stride, 4 * stride
    x = rng.standard_normal((B, C, H, W)).astype(np.float32)
    td = torch.float32 if dtype == "f32" else torch.bfloat16
    if dtype == "bf16":
        x = torch.from_numpy(x).to(torch.bfloat16).float().numpy()
    want = O.reorg(x, stride, 0)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu), dtype=td)
    yd, oc = ops.reorg(xd, C, stride, 0)
    got = ops.to_nchw(yd, oc).cpu().numpy()
    assert got.shape == want.shape and np.array_equal(got, want)
    dy = rng.standard_normal(want.shape).astype(np.float32)
    if dtype == "bf16":
        dy = torch.from_numpy(dy).to(torch.bfloat16).float().numpy()
    dyd = ops.to_nhwc(torch.from_numpy(dy).to(gpu), dtype=td)
    dxd, ic = ops.reorg_bwd(dyd, oc, stride, 0)
    dx = ops.to_nchw(dxd, ic).cpu().numpy()
    assert ic == C and np.array_equal(dx, O.reorg(dy, stride, 1))


@pytest.mark.parametrize("B,H,W,C", [(3, 19, 19, 512), (2, 13, 13, 48), (1, 20, 17, 16)])
def test_spp_maxpools_in_one_launch(gpu, B, H, W, C):
    """yc_maxpool_s1_multi (the SPP block's three stride-1 max-pools of one tensor, outputs = channel slices of the concat
    buffer) == three yc_maxpool_fwd calls, bit for bit (those are checked against the oracle above)."""
    import ctypes as C_
    from darknet_b200 import lib as L
    lib = L.lib()
    g = torch.Generator(device="cpu").manual_seed(H * 100 + C)
    cat = torch.randn((B, H, W, 4 * C), generator=g).to(torch.bfloat16).to(gpu)        # [mp13 | mp9 | mp5 | x] like route -1,-3,-5,-6
    x = L.view(cat, B, H, W, C, 4 * C, c_off=3 * C)
    sizes, pads = [5, 9, 13], [4, 8, 12]
    want = cat.clone()
    for o, (s, p) in enumerate(zip(sizes, pads)):
        y = L.view(want, B, H, W, C, 4 * C, c_off=(2 - o) * C)
        xw = L.view(want, B, H, W, C, 4 * C, c_off=3 * C)
        L.check(lib.yc_maxpool_fwd(xw, y, s, 1, p, 0, L.stream_ptr()), "yc_maxpool_fwd")
    got = cat.clone()
    xg = L.view(got, B, H, W, C, 4 * C, c_off=3 * C)
    ys = [L.view(got, B, H, W, C, 4 * C, c_off=(2 - o) * C) for o in range(3)]
    arr = (C_.POINTER(L.Tensor) * 3)(*[C_.pointer(y) for y in ys])
    assert all(lib.yc_maxpool_s1_multi_eligible(xg, y, s, 1, p) == 1 for y, s, p in zip(ys, sizes, pads))
    L.check(lib.yc_maxpool_s1_multi(xg, 3, arr, (C_.c_int * 3)(*sizes), (C_.c_int * 3)(*pads), L.stream_ptr()), "yc_maxpool_s1_multi")
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    assert not torch.equal(got[..., :C], cat[..., :C])


def test_spp_block_of_the_network_is_merged(gpu):
    """yolov4's SPP max-pools run as one launch under route elision (default) and as three without it: same outputs"""
    from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N
    text = cfgs.CONFIGS["yolov4"](w=320, h=320)
    weights = cfgmod.synth_weights(text)
    x = cfgmod.synth_images(2, 3, 320, 320, seed=9)
    outs, n_ops = [], []
    for flags in (L.NET_DEFAULT, L.NET_DEFAULT & ~L.NET_ELIDE_ROUTE):
        net = N.LoadSynthetic(text, weights, batch=2, flags=flags)
        pools = [i for i in range(net.n) if net.layer_type(i) == "maxpool"]
        assert len(pools) == 3
        N.NetworkPredict(net, x)
        n_ops.append(L.lib().yc_network_info(net.h, 6))
        outs.append([np.array(net.layer_output(i), copy=True) for i in pools + [pools[-1] + 1]])
        net.close()
    for a, b in zip(*outs):
        assert np.array_equal(a, b)
    routes = 21   # yolov4 has 21 [route] layers: at least one copy each without elision, and the pools stay apart
    assert n_ops[1] - n_ops[0] >= routes + 2, f"ops {n_ops}: the three pools are not one launch"
