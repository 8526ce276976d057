"""Parity of the training-mode conv operators (forward with batch statistics, BN backward, tcgen05 wgrad, dgrad
through the forward kernel) against the CPU oracle's restatement of ForwardConvolutionalLayer(train) /
BackwardConvolutionalLayer (pinned against the reference in tests/test_oracle_pin.py).

Inputs, weights and incoming deltas are rounded to bf16 first.  Tolerances (north_star bf16 bound): outputs and
gradients within 1e-2 of the tensor's max (2e-2 for dX / dW behind a batchnorm backward, whose cancellation
amplifies the bf16 rounding of z and dz), statistics within 1e-3.
"""
import numpy as np
import pytest
import torch

from oracle.pyoracle import O

pytestmark = pytest.mark.gpu


def _r(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).float().numpy()


def _relmax(got, ref):
    return float(np.abs(got - ref).max() / (np.abs(ref).max() + 1e-12))


def _rell2(got, ref):
    return float(np.linalg.norm((got - ref).ravel()) / (np.linalg.norm(ref.ravel()) + 1e-30))


def _case(gpu, B, C, H, W, N, k, stride, act, bn=True, seed=0, accumulate=False, pre_dw=False):
    from darknet_b200 import ops
    rng = np.random.default_rng(seed)
    pad = k // 2
    x = _r(rng.standard_normal((B, C, H, W)))
    w = _r(rng.standard_normal((N, C, k, k)) * np.sqrt(2.0 / (C * k * k)))
    if bn:
        biases = rng.uniform(-0.5, 0.5, N).astype(np.float32)
        scales = rng.uniform(0.5, 1.5, N).astype(np.float32)
    else:
        biases = rng.uniform(-0.5, 0.5, N).astype(np.float32)
        scales = None
    fwd = O.conv_forward_train(x, w, biases, k, stride, pad, act, scales=scales)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
    conv = ops.ConvTrain(xd, (B, H, W, C), torch.from_numpy(w), torch.from_numpy(biases), k, stride, pad, act,
                         scales=None if scales is None else torch.from_numpy(scales))
    y = conv.forward()
    torch.cuda.synchronize()
    got = ops.to_nchw(y, N).cpu().numpy()
    e_fwd = _relmax(got, fwd["out"])
    assert e_fwd < 1e-2, f"forward {e_fwd:.3e}"
    if bn:
        assert _relmax(conv.mean.cpu().numpy(), fwd["mean"]) < 2e-3 or np.abs(conv.mean.cpu().numpy() - fwd["mean"]).max() < 2e-3
        assert _relmax(conv.var.cpu().numpy(), fwd["var"]) < 2e-3
        assert _relmax(conv.rvar.cpu().numpy(), fwd["rolling_var"]) < 2e-3
    # backward
    oh, ow = fwd["out"].shape[2:]
    delta = _r(rng.standard_normal((B, N, oh, ow)) * 0.1)
    wu0 = None
    if pre_dw:
        wu0 = rng.standard_normal(w.shape).astype(np.float32)
        L = __import__("darknet_b200.lib", fromlist=["x"])
        t = torch.from_numpy(wu0).to(gpu)
        L.check(L.lib().yc_conv_wgrad_pack(t.data_ptr(), N, C, k, conv.dw_packed.data_ptr(), L.stream_ptr()), "pack")
    # the backward operators are checked on the SAME saved forward state the GPU holds (bf16 z, its statistics):
    # with the oracle's own fp32 z, elements whose pre-activation is within bf16 rounding of 0 flip the leaky
    # branch, which is forward rounding noise and not a property of the backward kernels.
    fwd2 = dict(fwd)
    if bn:
        zg = ops.to_nchw(conv.z, N).cpu().numpy()
        mg, vg = conv.mean.cpu().numpy(), conv.var.cpu().numpy()
        xn = (zg - mg[None, :, None, None]) / np.sqrt(vg[None, :, None, None] + np.float32(1e-6))
        pre = (xn * scales[None, :, None, None] + biases[None, :, None, None]).astype(np.float32)
        fwd2.update(z=zg, mean=mg, var=vg, xnorm=xn.astype(np.float32), pre=pre, out=O.activate(pre, act))
    else:
        fwd2.update(out=got, pre=got)
    ref = O.conv_backward(x, w, fwd2, delta, k, stride, pad, act, scales=scales, weight_updates=wu0)
    dyd = ops.to_nhwc(torch.from_numpy(delta).to(gpu))
    if accumulate:
        dx0 = _r(rng.standard_normal(x.shape) * 0.05)
        conv.dx.copy_(ops.to_nhwc(torch.from_numpy(dx0).to(gpu)))
    dx = conv.backward(dyd, accumulate=accumulate)
    torch.cuda.synchronize()
    tol = 2e-2 if bn else 1e-2
    got_dz = ops.to_nchw(conv.dz, N).cpu().numpy()
    assert _relmax(got_dz, ref["delta"]) < tol, f"dz {_relmax(got_dz, ref['delta']):.3e}"
    assert _rell2(got_dz, ref["delta"]) < 1e-2, f"dz rel-L2 {_rell2(got_dz, ref['delta']):.3e}"
    if conv.cd > N:
        assert float(conv.dz[..., N:].abs().max()) == 0.0, "pad lanes of dz must be zero"
    got_dw = conv.weight_updates().cpu().numpy()
    e_dw = _relmax(got_dw, ref["weight_updates"])
    assert e_dw < tol, f"weight_updates {e_dw:.3e} {conv.winfo()}"
    assert _rell2(got_dw, ref["weight_updates"]) < 1e-2, f"weight_updates rel-L2 {_rell2(got_dw, ref['weight_updates']):.3e}"
    if bn:
        e_su = _relmax(conv.scale_updates.cpu().numpy(), ref["scale_updates"])
        assert e_su < tol, f"scale_updates {e_su:.3e}"
    else:
        e_bu = _relmax(conv.bias_updates.cpu().numpy(), ref["bias_updates"])
        assert e_bu < tol, f"bias_updates {e_bu:.3e}"
    got_dx = ops.to_nchw(dx, C).cpu().numpy()
    want_dx = ref["dx"] + (dx0 if accumulate else 0)
    e_dx = _relmax(got_dx, want_dx)
    assert e_dx < tol, f"dx {e_dx:.3e}"
    assert _rell2(got_dx, want_dx) < 1e-2, f"dx rel-L2 {_rell2(got_dx, want_dx):.3e}"
    return e_fwd, e_dw, e_dx


@pytest.mark.parametrize("C,N,HW,k", [(64, 64, 16, 3), (128, 128, 13, 3), (64, 128, 19, 1), (128, 64, 16, 1),
                                      (256, 512, 8, 3), (512, 256, 8, 1)])
def test_train_bn_leaky(gpu, C, N, HW, k):
    _case(gpu, 2, C, HW, HW, N, k, 1, "leaky")


@pytest.mark.parametrize("act", ["mish", "swish", "logistic", "linear", "relu"])
def test_train_bn_activations(gpu, act):
    _case(gpu, 2, 64, 12, 12, 64, 3, 1, act)


def test_train_stride2(gpu):
    _case(gpu, 2, 64, 16, 16, 128, 3, 2, "leaky")
    _case(gpu, 2, 32, 17, 23, 64, 3, 2, "leaky")


def test_train_small_and_odd_channels(gpu):
    _case(gpu, 2, 3, 32, 32, 32, 3, 1, "leaky")       # stem
    _case(gpu, 2, 32, 16, 16, 64, 3, 1, "leaky")
    _case(gpu, 2, 96, 10, 10, 192, 3, 1, "leaky")
    _case(gpu, 2, 384, 7, 7, 96, 1, 1, "leaky")


def test_train_head_no_bn(gpu):
    _case(gpu, 2, 128, 13, 13, 255, 1, 1, "linear", bn=False)
    _case(gpu, 2, 64, 8, 8, 18, 1, 1, "linear", bn=False)
    _case(gpu, 2, 64, 8, 8, 64, 3, 1, "leaky", bn=False)


def test_train_accumulate_semantics(gpu):
    _case(gpu, 2, 64, 12, 12, 64, 3, 1, "leaky", accumulate=True)   # dgrad += (cuDNN-like), B6
    _case(gpu, 2, 64, 12, 12, 64, 3, 1, "leaky", pre_dw=True)       # wgrad += existing weight_updates, B7


def test_train_many_pixels_split_k(gpu):
    _case(gpu, 8, 64, 52, 52, 64, 3, 1, "leaky")
    _case(gpu, 4, 128, 38, 38, 256, 3, 1, "leaky")


def test_wgrad_tap_groups_and_single_tap_units(gpu):
    """wgrad work units: a 3x3 layer's unit covers a kernel row (or the whole filter) and shares the dY boxes, with one
    TMEM accumulator per tap; yc_conv_wgrad_set_tap_groups(0) plans one tap per unit as in round 1.  Both schedules on the
    channel counts that select each variant: C = 32 / 64 (row groups, double-buffered), 128 (row groups, one accumulator
    set), 256 (always one tap per unit)."""
    from darknet_b200 import lib as L
    lib = L.lib()
    cases = [(2, 32, 16, 16, 64), (4, 64, 26, 26, 64), (2, 128, 13, 13, 128), (2, 256, 8, 8, 256)]
    try:
        for on in (1, 0):
            lib.yc_conv_wgrad_set_tap_groups(on)
            for (B, C, H, W, N) in cases:
                _case(gpu, B, C, H, W, N, 3, 1, "leaky", seed=on)
    finally:
        lib.yc_conv_wgrad_set_tap_groups(1)
