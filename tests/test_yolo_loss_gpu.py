"""Parity of the on-device YOLO loss / delta kernels against the UNMODIFIED reference's ForwardYoloLayer (train)
(yolo_layer.cpp:409-772), on the reference's own decoded head: same inputs, so deltas must agree to fp32 rounding
(1e-5 absolute on values of order 1) and the cost to 1e-4 relative (the reference sums 1e5 squares sequentially in
float, utils.cpp:271-280)."""
import os

import numpy as np
import pytest
import torch

from darknet_b200 import cfg as cfgmod, cfgs
from oracle import pyoracle
from tests import netutil

pytestmark = pytest.mark.gpu

ANCH = [4, 5, 8, 10, 13, 16, 20, 24, 30, 35, 40, 50]


def _ref_heads(text, tag, B, classes, seed):
    cpath = os.path.join(netutil._TMP, tag + ".cfg")
    wpath = os.path.join(netutil._TMP, tag + ".weights")
    with open(cpath, "w") as f:
        f.write(text)
    weights = cfgmod.synth_weights(text, seed=42)
    cfgmod.write_weights(wpath, text, weights)
    ref = pyoracle.RefNet(cpath, wpath, train=True)
    x = cfgmod.synth_images(B, 3, 64, 64, seed=seed)
    truth = cfgmod.synth_truth(B, 90, classes=classes, seed=seed + 1)
    ref.forward(x, truth, train=1)
    heads = []
    for i in range(ref.n):
        if ref.layer_int(i, "classes") > 0 and ref.layer_buf(i, "cost") is not None and ref.layer_int(i, "n") == 3 and ref.layer_int(i, "total") == 6:
            out = ref.layer_output(i)
            delta = ref.layer_buf(i, "delta").reshape(out.shape)
            heads.append((i, out, delta, float(ref.layer_buf(i, "cost")[0])))
    ref.close()
    return truth, heads


@pytest.mark.parametrize("iou_loss,extra", [("ciou", {}), ("mse", {}), ("giou", {}), ("diou", {}), ("iou", {}),
                                            ("ciou", {"focal_loss": 1}), ("ciou", {"label_smooth_eps": 0.1}),
                                            ("ciou", {"max_delta": 2}), ("mse", {"iou_thresh": 1})])
def test_yolo_loss_vs_reference(gpu, iou_loss, extra):
    from darknet_b200 import ops
    pytest.importorskip("numpy")
    if not pyoracle.have_ref():
        pytest.skip("oracle/_ref not built")
    B, classes = 4, 3
    text = cfgs.micro(batch=B, subdivisions=1, classes=classes).replace("iou_loss=ciou", "iou_loss=" + iou_loss)
    kw = dict(iou_thresh=0.213, iou_normalizer=0.07, cls_normalizer=1.0, ignore_thresh=.7, truth_thresh=1.0)
    for k, v in extra.items():
        if k == "iou_thresh":
            text = text.replace("iou_thresh=0.213", f"iou_thresh={v}")
        else:
            text = text.replace("random=0", f"random=0\n{k}={v}")
        kw[k] = v
    tag = "loss_" + iou_loss + "_" + "_".join(f"{k}{v}" for k, v in extra.items())
    truth, heads = _ref_heads(text, tag, B, classes, seed=11)
    assert len(heads) == 2
    masks = {17: [3, 4, 5], 24: [0, 1, 2]}
    for (i, out, dref, cost) in heads:
        o = torch.from_numpy(np.ascontiguousarray(out)).to(gpu)
        prev = torch.zeros((B, out.shape[2], out.shape[3], 32), dtype=torch.bfloat16, device=gpu)
        delta, stats = ops.yolo_loss(o, truth, ANCH, masks[i], classes, 64, 64, iou_loss=iou_loss, prev_delta=prev, **kw)
        torch.cuda.synchronize()
        d = delta.cpu().numpy()
        err = np.abs(d - dref).max()
        assert err < 2e-5, f"head {i}: delta max abs err {err:.3e} (max |delta| {np.abs(dref).max():.3f})"
        st = stats.cpu().numpy()
        assert abs(st[0] - cost) <= 1e-4 * abs(cost) + 1e-5, f"head {i}: cost {st[0]} vs {cost}"
        # BackwardYoloLayer: prev.delta (zero here) += delta, written as NHWC bf16
        pv = prev.float().cpu().numpy()[..., :out.shape[1]].transpose(0, 3, 1, 2)
        assert np.abs(pv - dref).max() <= 8e-3 * max(1.0, np.abs(dref).max())


@pytest.mark.parametrize("iou_loss,extra", [("giou", {}), ("mse", {}), ("ciou", {}), ("diou", {}), ("giou", {"label_smooth_eps": 0.1}),
                                            ("mse", {"max_delta": 2}), ("mse", {"uc_normalizer": 0.5}),
                                            ("giou", {"iou_thresh": 1})])
def test_gaussian_yolo_loss_vs_reference(gpu, iou_loss, extra):
    """[Gaussian_yolo] (BASELINE configs[4]): ForwardGaussianYoloLayer's training half (gaussian_yolo_layer.cpp:525-857)
    on the reference's own decoded head.  The reference evaluates the Gaussian terms in mixed float/double and is built
    with -Ofast, so deltas agree to 5e-5 absolute rather than 2e-5."""
    from darknet_b200 import ops
    if not pyoracle.have_ref():
        pytest.skip("oracle/_ref not built")
    B, classes = 4, 3
    text = cfgs.micro(batch=B, subdivisions=1, classes=classes, gaussian=True).replace("iou_loss=giou", "iou_loss=" + iou_loss)
    kw = dict(iou_thresh=0.213, iou_normalizer=0.5, cls_normalizer=1.0, ignore_thresh=.7, truth_thresh=1.0, uc_normalizer=1.0)
    for k, v in extra.items():
        if k == "iou_thresh":
            text = text.replace("iou_thresh=0.213", f"iou_thresh={v}")
        elif k == "uc_normalizer":
            text = text.replace("uc_normalizer=1.0", f"uc_normalizer={v}")
        else:
            text = text.replace("random=0", f"random=0\n{k}={v}")
        kw[k] = v
    tag = "gloss_" + iou_loss + "_" + "_".join(f"{k}{v}" for k, v in extra.items())
    truth, heads = _ref_heads(text, tag, B, classes, seed=21)
    assert len(heads) == 2
    masks = {17: [3, 4, 5], 24: [0, 1, 2]}
    for (i, out, dref, cost) in heads:
        assert out.shape[1] == 3 * (9 + classes)
        o = torch.from_numpy(np.ascontiguousarray(out)).to(gpu)
        prev = torch.zeros((B, out.shape[2], out.shape[3], 40), dtype=torch.bfloat16, device=gpu)
        delta, stats = ops.yolo_loss(o, truth, ANCH, masks[i], classes, 64, 64, iou_loss=iou_loss, prev_delta=prev,
                                     gaussian=True, **kw)
        torch.cuda.synchronize()
        d = delta.cpu().numpy()
        assert np.abs(dref).max() > 0.1, "reference produced no deltas: the case tests nothing"
        err = np.abs(d - dref).max()
        assert err < 5e-5 * max(1.0, np.abs(dref).max()), f"head {i}: delta max abs err {err:.3e} (max |delta| {np.abs(dref).max():.3f})"
        st = stats.cpu().numpy()
        assert abs(st[0] - cost) <= 1e-4 * abs(cost) + 1e-5, f"head {i}: cost {st[0]} vs {cost}"
        pv = prev.float().cpu().numpy()[..., :out.shape[1]].transpose(0, 3, 1, 2)
        assert np.abs(pv - dref).max() <= 8e-3 * max(1.0, np.abs(dref).max())
