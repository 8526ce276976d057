"""Pins the oracle (oracle/yc_oracle.c restatement + oracle/pynet.py walker) against the UNMODIFIED reference CPU
path compiled from /root/reference (oracle/_ref/libyolo_ref.so).  CPU only.

The reference ships no golden vectors, so the pin is the reference itself run here; committed
fixtures of those runs live in tests/golden/ (tests/test_golden.py)."""
import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod
from oracle import pyoracle
from oracle.pyoracle import O
from tests.netutil import materialise

pytestmark = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built")


def test_gemm_matches_reference(built):
    rng = np.random.default_rng(0)
    L = pyoracle.ref_lib()
    for (TA, TB, M, N, K) in [(0, 0, 64, 75, 27), (0, 0, 32, 361, 288), (0, 1, 16, 27, 100), (1, 0, 27, 100, 16)]:
        A = rng.standard_normal((K, M) if TA else (M, K)).astype(np.float32)
        B = rng.standard_normal((N, K) if TB else (K, N)).astype(np.float32)
        C0 = rng.standard_normal((M, N)).astype(np.float32)
        cr, co = C0.copy(), C0.copy()
        L.ref_gemm(TA, TB, M, N, K, 1.0, A.reshape(-1), A.shape[1], B.reshape(-1), B.shape[1], 1.0, cr.reshape(-1), N)
        pyoracle.oracle_lib().yo_gemm(TA, TB, M, N, K, 1.0, A.reshape(-1), A.shape[1], B.reshape(-1), B.shape[1], 1.0,
                                      co.reshape(-1), N)
        np.testing.assert_allclose(co, cr, rtol=2e-5, atol=2e-5)


def test_im2col_matches_reference(built):
    rng = np.random.default_rng(1)
    for (c, h, w, k, pad, stride, dil) in [(3, 9, 11, 3, 1, 1, 1), (8, 10, 10, 3, 1, 2, 1), (4, 7, 7, 1, 0, 1, 1), (2, 12, 9, 3, 2, 1, 2)]:
        im = rng.standard_normal((c, h, w)).astype(np.float32)
        oh = (h + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
        ow = (w + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
        a = np.zeros(c * k * k * oh * ow, np.float32)
        b = np.zeros_like(a)
        pyoracle.ref_lib().ref_im2col(im.reshape(-1), c, h, w, k, pad, stride, dil, a)
        pyoracle.oracle_lib().yo_im2col(im.reshape(-1), c, h, w, k, pad, stride, dil, b)
        assert np.array_equal(a, b)


@pytest.mark.parametrize("act", ["linear", "leaky", "mish", "logistic", "swish", "relu"])
def test_activations_match_reference(built, act):
    x = np.concatenate([np.linspace(-30, 30, 2001), [-0.0, 0.0, 1e-8, -1e-8, 20.0, -20.0, 20.0001, -20.0001]]).astype(np.float32)
    L = pyoracle.ref_lib()
    out = np.zeros_like(x)
    saved = np.zeros_like(x)
    L.ref_activate(x.copy(), x.size, L.ref_activation_id(act.encode()), saved, out)
    got = O.activate(x, act)
    # mish/tanh: 2/(1+expf(-2*sp))-1 cancels near sp->0, so libm-vs-vectorised expf ulps show up as ~6e-8*|x| absolute
    np.testing.assert_allclose(got, out, rtol=3e-6, atol=3e-6)


@pytest.mark.parametrize("name,kw", [("micro", {}), ("micro", {"act": "mish"}), ("micro", {"gaussian": True})])
def test_network_walk_matches_reference(built, name, kw):
    """Every layer of the micro net: C restatement == reference ForwardNetwork (fused BN, inference)."""
    from oracle import pynet
    text, cpath, wpath = materialise(name, **kw)
    weights, _ = cfgmod.read_weights(wpath, text)
    x = cfgmod.synth_images(1, 3, 64, 64)
    ref = pyoracle.RefNet(cpath, wpath, train=False, fuse=True)
    ref.predict(x)
    outs = pynet.forward(text, weights, x, fuse_bn=True, infer_avx_maxpool=True)
    for i, o in enumerate(outs):
        r = ref.layer_output(i)
        assert o.shape == r.shape
        err = np.abs(o - r).max() / (np.abs(r).max() + 1e-12)
        assert err < 2e-5, f"layer {i}: {err}"  # -Ofast reassociation in the reference build
    ref.close()


def test_detections_and_nms_match_reference(built):
    text, cpath, wpath = materialise("micro", head_bias=-3.0, head_gain=40.0)
    x = cfgmod.synth_images(1, 3, 64, 64)
    ref = pyoracle.RefNet(cpath, wpath, train=False, fuse=True)
    ref.predict(x)
    _, layers = cfgmod.walk_shapes(text)
    rows = []
    for L in layers:
        if L.kind != "yolo":
            continue
        mask = [int(v) for v in L.opts["mask"].split(",")]
        anchors = [float(v) for v in L.opts["anchors"].split(",")]
        pred = ref.layer_output(L.index)
        rows.append(O.yolo_detections(pred, 0, len(mask), int(L.opts["classes"]), anchors, mask, 64, 64, 0.25))
    mine = np.concatenate(rows, axis=0)
    r7, rprobs = ref.boxes(0.25, 0.0)
    assert mine.shape[0] == r7.shape[0] > 0
    np.testing.assert_allclose(mine[:, :5], r7[:, :5], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(mine[:, 5:], rprobs, rtol=1e-6, atol=1e-7)
    # NMS: the restatement uses a stable sort, so compare the surviving set, not the array order
    post = O.nms_sort(mine, 3, 0.45)
    r7n, rprobsn = ref.boxes(0.25, 0.45)
    a = sorted((tuple(np.round(post[i, :4], 3)), int(c)) for i in range(post.shape[0]) for c in np.nonzero(post[i, 5:] > 0)[0])
    b = sorted((tuple(np.round(r7n[i, :4], 3)), int(c)) for i in range(r7n.shape[0]) for c in np.nonzero(rprobsn[i] > 0)[0])
    assert a == b
    ref.close()



@pytest.mark.parametrize("act", ["linear", "leaky", "logistic"])
def test_scale_channels_restatement_vs_reference(act):
    """ForwardScaleChannelsLayer / BackwardScaleChannelsLayer (scale_channels_layer.c:72-131) — the reference's own
    functions on a hand-built two-layer network vs oracle/yc_oracle.c (bit for bit where no libm call is involved)."""
    import ctypes as C
    rng = np.random.default_rng(3)
    B, Cc, H, W = 2, 24, 5, 7
    from_out = rng.standard_normal((B, Cc, H, W)).astype(np.float32)
    scale_in = rng.uniform(0.1, 2.0, (B, Cc)).astype(np.float32)
    delta = rng.standard_normal((B, Cc, H, W)).astype(np.float32)
    ds0 = rng.standard_normal((B, Cc)).astype(np.float32)     # the gradients ACCUMULATE onto what is there
    df0 = rng.standard_normal((B, Cc, H, W)).astype(np.float32)
    out = O.scale_channels(from_out, scale_in, act)
    d, ds, df = O.scale_channels_backward(delta, out, from_out, scale_in, act, ds0, df0)
    r_out = np.zeros_like(out)
    r_d, r_ds, r_df = delta.copy(), ds0.copy(), df0.copy()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    pyoracle.ref_lib().ref_scale_channels(from_out.reshape(-1), scale_in.reshape(-1), B, Cc, H, W, pyoracle.ACT[act],
                                          r_out.reshape(-1), vp(r_d), vp(r_ds), vp(r_df))
    if act == "logistic":   # the reference is built with -Ofast: its expf differs from libm's in the last bits
        np.testing.assert_allclose(out, r_out, rtol=2e-6, atol=1e-6)
        np.testing.assert_allclose(d, r_d, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(df, r_df, rtol=1e-5, atol=1e-6)
    else:
        assert np.array_equal(out, r_out)
        assert np.array_equal(d, r_d) and np.array_equal(df, r_df)
    np.testing.assert_allclose(ds, r_ds, rtol=1e-5, atol=1e-5)   # -Ofast may reassociate the channel sums



@pytest.mark.parametrize("stride,forward", [(2, 0), (2, 1), (4, 0)])
def test_reorg_restatement_vs_reference(stride, forward):
    rng = np.random.default_rng(4)
    B, Cc, H, W = 2, 3 * (stride * stride if forward else 1), 8, 8
    x = rng.standard_normal((B, Cc, H, W)).astype(np.float32)
    got = O.reorg(x, stride, forward)
    s2 = stride * stride
    oc, oh, ow = (Cc, H, W) if forward else (Cc * s2, H // stride, W // stride)
    want = np.zeros_like(got)
    pyoracle.ref_lib().ref_reorg(x.reshape(-1), ow, oh, oc, B, stride, forward, want.reshape(-1))
    assert np.array_equal(got, want)
    if not forward:  # BackwardReorgLayer = the inverse rearrangement: round trip is the identity
        assert np.array_equal(O.reorg(got, stride, 1), x)
