"""Pin of the oracle's TRAINING restatement (oracle/yc_oracle.c: yo_conv_forward_train / yo_conv_backward) against the
UNMODIFIED reference CPU path (oracle/_ref): a 3-layer net [conv+BN+act, stride s] -> [1x1 conv, linear, bias] ->
[yolo], ForwardNetwork(train) + BackwardNetwork.  The reference's own YOLO delta is fed to the restatement, so every
compared quantity depends only on the conv / batchnorm code under test."""
import os
import tempfile

import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod
from oracle import pyoracle
from oracle.pyoracle import O

pytestmark = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built (needs /root/reference)")

CFG = """[net]
batch={B}
subdivisions=1
width=16
height=16
channels=3
momentum=0.9
decay=0.0005
learning_rate=0.001

[convolutional]
batch_normalize=1
filters=8
size=3
stride={s}
pad=1
activation={act}

[convolutional]
filters=24
size=1
stride=1
pad=1
activation=linear

[yolo]
mask=0,1,2
anchors=2,3, 4,5, 6,8
classes=3
num=3
ignore_thresh=.7
truth_thresh=1
iou_loss=ciou
iou_normalizer=0.07
"""


@pytest.mark.parametrize("act,s", [("leaky", 1), ("mish", 1), ("leaky", 2), ("swish", 1), ("logistic", 2)])
def test_conv_train_restatement_vs_reference(act, s):
    B = 3
    text = CFG.format(B=B, act=act, s=s)
    d = tempfile.mkdtemp(prefix="yc_pin_")
    cpath, wpath = os.path.join(d, "n.cfg"), os.path.join(d, "n.weights")
    open(cpath, "w").write(text)
    weights = cfgmod.synth_weights(text, seed=3)
    cfgmod.write_weights(wpath, text, weights)
    ref = pyoracle.RefNet(cpath, wpath, train=True)
    x = cfgmod.synth_images(B, 3, 16, 16, seed=9)
    truth = cfgmod.synth_truth(B, 90, classes=3, seed=10)
    ref.forward_backward(x, truth)
    w0, w1 = weights[0], weights[1]
    W0 = w0["weights"].reshape(8, 3, 3, 3)
    W1 = w1["weights"].reshape(24, 8, 1, 1)
    f0 = O.conv_forward_train(x, W0, w0["biases"], 3, s, 1, act, scales=w0["scales"],
                              rolling=(w0["rolling_mean"], w0["rolling_variance"]))
    np.testing.assert_allclose(f0["out"], ref.layer_output(0), rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(f0["mean"], ref.layer_buf(0, "mean"), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(f0["var"], ref.layer_buf(0, "variance"), rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(f0["rolling_mean"], ref.layer_buf(0, "rolling_mean"), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(f0["rolling_var"], ref.layer_buf(0, "rolling_variance"), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(f0["z"].ravel(), ref.layer_buf(0, "x"), rtol=2e-5, atol=2e-6)
    f1 = O.conv_forward_train(f0["out"], W1, w1["biases"], 1, 1, 0, "linear")
    np.testing.assert_allclose(f1["out"], ref.layer_output(1), rtol=2e-5, atol=2e-6)
    # backward, fed with the reference's own YOLO delta (= layer 1's delta: linear, no BN -> unchanged by its backward)
    ydelta = ref.layer_buf(2, "delta").reshape(f1["out"].shape)
    np.testing.assert_allclose(ref.layer_buf(1, "delta").reshape(ydelta.shape), ydelta, rtol=0, atol=0)
    r1 = O.conv_backward(f0["out"], W1, f1, ydelta, 1, 1, 0, "linear")
    np.testing.assert_allclose(r1["weight_updates"].ravel(), ref.layer_buf(1, "weight_updates"), rtol=2e-4, atol=2e-6)
    np.testing.assert_allclose(r1["bias_updates"], ref.layer_buf(1, "bias_updates"), rtol=2e-4, atol=2e-6)
    r0 = O.conv_backward(x, W0, f0, r1["dx"], 3, s, 1, act, scales=w0["scales"], want_dx=False)
    d0 = ref.layer_buf(0, "delta").reshape(r0["delta"].shape)
    scale = np.abs(d0).max()
    assert np.abs(r0["delta"] - d0).max() <= 2e-4 * scale, "layer-0 delta after the batchnorm backward"
    wu = ref.layer_buf(0, "weight_updates")
    assert np.abs(r0["weight_updates"].ravel() - wu).max() <= 2e-4 * np.abs(wu).max()
    su = ref.layer_buf(0, "scale_updates")
    assert np.abs(r0["scale_updates"] - su).max() <= 2e-4 * np.abs(su).max() + 1e-6
    assert np.abs(ref.layer_buf(0, "bias_updates")).max() == 0.0   # the BN shift gets no gradient on the CPU path
    ref.close()


@pytest.mark.parametrize("act", ["mish", "leaky"])
def test_training_walker_fp32_vs_reference(act):
    """oracle/pytrain.py in fp32 mode (whole micro net, two steps with an SGD update in between) == the unmodified
    reference.  mish is smooth: everything to 1e-4.  leaky' jumps at 0, so a handful of elements whose pre-activation
    is ~1e-6 flip under a different summation order even in fp32: gradients to 8e-2, everything else tight."""
    from oracle import pytrain
    from tests import netutil
    from tests.netutil import rel_l2
    B = 4
    text, cpath, wpath = netutil.materialise("micro", batch=B, subdivisions=1, act=act)
    ref = pyoracle.RefNet(cpath, wpath, train=True)
    W, _ = cfgmod.read_weights(wpath, text)
    st = pytrain.Step(text, W)
    tol_g = 1e-4 if act == "mish" else 8e-2
    for step in range(2):
        x = cfgmod.synth_images(B, 3, 64, 64, seed=5 + step)
        truth = cfgmod.synth_truth(B, 90, classes=3, seed=6 + step)
        cr = ref.forward_backward(x, truth)
        c = st.forward_backward(x, truth)
        assert abs(c - cr) <= 1e-4 * abs(cr)
        for L in st.layers:
            i = L.index
            assert rel_l2(st.out[i], ref.layer_output(i)) < (1e-4 if act == "mish" else 2e-3)
            d = ref.layer_buf(i, "delta")
            if d is not None:
                assert rel_l2(st.delta[i], d.reshape(st.delta[i].shape)) < tol_g, (i, L.kind)
            if i in st.P:
                assert rel_l2(st.P[i]["weight_updates"].ravel(), ref.layer_buf(i, "weight_updates")) < tol_g
                if L.bn:
                    assert rel_l2(st.P[i]["scale_updates"], ref.layer_buf(i, "scale_updates")) < tol_g
                    assert rel_l2(st.saved[i]["var"], ref.layer_buf(i, "variance")) < (1e-4 if act == "mish" else 2e-3)
                    assert rel_l2(st.P[i]["rolling_variance"], ref.layer_buf(i, "rolling_variance")) < (1e-4 if act == "mish" else 2e-3)
                else:
                    assert rel_l2(st.P[i]["bias_updates"], ref.layer_buf(i, "bias_updates")) < 1e-4
        ref.L.ref_update(ref.h)
        st.update()
        for i in st.P:
            assert rel_l2(st.P[i]["weights"].ravel(), ref.layer_buf(i, "weights")) < (1e-5 if act == "mish" else 2e-3)
    ref.close()


@pytest.mark.parametrize("iou_loss", ["ciou", "mse", "giou", "diou"])
def test_yolo_loss_restatement_vs_reference(iou_loss):
    """oracle/yc_oracle_loss.c == the reference's ForwardYoloLayer(train) on the reference's own decoded heads."""
    from darknet_b200 import cfgs
    from oracle import pytrain
    from tests import netutil
    B, classes = 4, 3
    text = cfgs.micro(batch=B, subdivisions=1, classes=classes).replace("iou_loss=ciou", "iou_loss=" + iou_loss)
    d = tempfile.mkdtemp(prefix="yc_pin_")
    cpath, wpath = os.path.join(d, "n.cfg"), os.path.join(d, "n.weights")
    open(cpath, "w").write(text)
    cfgmod.write_weights(wpath, text, cfgmod.synth_weights(text, seed=42))
    ref = pyoracle.RefNet(cpath, wpath, train=True)
    x = cfgmod.synth_images(B, 3, 64, 64, seed=11)
    truth = cfgmod.synth_truth(B, 90, classes=classes, seed=12)
    ref.forward(x, truth, train=1)
    _, layers = cfgmod.walk_shapes(text)
    for L in layers:
        if L.kind != "yolo":
            continue
        out = ref.layer_output(L.index)
        mask = [int(v) for v in L.opts["mask"].split(",")]
        anchors = [float(v) for v in L.opts["anchors"].split(",")]
        delta, cost = pytrain.yolo_loss(out, truth, anchors, mask, L.opts, 64, 64)
        dref = ref.layer_buf(L.index, "delta").reshape(out.shape)
        assert np.abs(delta - dref).max() < 2e-5
        cref = float(ref.layer_buf(L.index, "cost")[0])
        assert abs(cost - cref) <= 1e-4 * abs(cref) + 1e-5
    ref.close()
