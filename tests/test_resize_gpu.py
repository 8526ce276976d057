"""yc_network_resize (ResizeNetwork, network.cpp:255-383; detector.cpp:150-182 calls it every ten iterations under
random=1): a resized network computes what a network LOADED at that size computes (same bits: same plans, same weights),
and a training network keeps its parameters, rolling statistics, momentum and counters across a resize."""
import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod, cfgs

pytestmark = pytest.mark.gpu


def _heads(net, N, x):
    N.NetworkPredict(net, x)
    return [N.yolo_output(net, i).copy() for i in net.yolo_layers()]


@pytest.mark.parametrize("name,B,sizes", [("yolov4-tiny", 4, [(416, 416), (320, 320), (608, 352), (416, 416)]),
                                          ("yolov4", 2, [(608, 608), (416, 416), (608, 608)])])
def test_resized_inference_network_equals_one_loaded_at_that_size(gpu, name, B, sizes):
    from darknet_b200 import lib as L, network as N
    w0, h0 = sizes[0]
    text0 = cfgs.CONFIGS[name](w=w0, h=h0, batch=B)
    weights = cfgmod.synth_weights(text0)
    net = N.LoadSynthetic(text0, weights, batch=B)
    for (w, h) in sizes:
        N.ResizeNetwork(net, w, h)
        assert (net.w, net.hh) == (w, h)
        fresh = N.LoadSynthetic(cfgs.CONFIGS[name](w=w, h=h, batch=B), weights, batch=B)
        x = cfgmod.synth_images(B, 3, h, w, seed=w + h)
        a, b = _heads(net, N, x), _heads(fresh, N, x)
        assert len(a) == len(b) > 0
        for u, v in zip(a, b):
            assert u.shape == v.shape and np.array_equal(u, v), f"{name} {w}x{h}: resized != loaded at that size"
        # detections through the resized network's workspace
        # (yolov4 with random heads: every cell of the 76x76 head is a candidate, which also takes yc_network_detect through
        # its regrow path; the host NMS of 2 x 54k boxes would take minutes, so it is off there)
        nms = 0.45 if name == "yolov4-tiny" else 0.0
        da, db = N.DetectBatch(net, 0.25, nms), N.DetectBatch(fresh, 0.25, nms)
        assert np.array_equal(da[0], db[0]) and np.array_equal(da[1], db[1])
        fresh.close()
    net.close()


def _snapshot(net, N, out, tag):
    for i in range(net.n):
        if net.layer_info(i, 0) == 0:
            out[f"{tag}.L{i}.weights"] = N.layer_param(net, i, "weights")
            out[f"{tag}.L{i}.weight_updates"] = N.layer_param(net, i, "weight_updates")
            out[f"{tag}.L{i}.biases"] = N.layer_param(net, i, "biases")
            if net.layer_info(i, 9):
                out[f"{tag}.L{i}.rolling_mean"] = N.layer_param(net, i, "rolling_mean")
                out[f"{tag}.L{i}.rolling_variance"] = N.layer_param(net, i, "rolling_variance")


def _train_run(detour):
    from darknet_b200 import lib as L, network as N
    B, hw, classes = 4, 416, 80
    text = cfgs.CONFIGS["yolov4-tiny"](batch=B, subdivisions=1, w=hw, h=hw)
    net = N.LoadSynthetic(text, cfgmod.synth_weights(text), batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    out = {}
    for s in range(3):
        if detour and s == 1:
            N.ResizeNetwork(net, 320, 352)   # there ...
            assert net.layer_info(net.n - 1, 2) == 352 // 16 and net.layer_info(net.n - 1, 3) == 320 // 16
            N.ResizeNetwork(net, hw, hw)     # ... and back, between an update and the next forward
        x = cfgmod.synth_images(B, 3, hw, hw, seed=5 + s)
        truth = cfgmod.synth_truth(B, 90, classes=classes, seed=6 + s)
        out[f"cost{s}"] = np.float32(N.TrainNetworkDatum(net, x, truth))
        N.UpdateNetwork(net)
        _snapshot(net, N, out, f"s{s}")
        out[f"seen{s}"] = np.int64(net.info(10))
    net.close()
    return out


def _same(a, b, what):
    assert a.keys() == b.keys()
    bad = [k for k in a if not np.array_equal(a[k], b[k])]
    assert not bad, f"{what}: {len(bad)} of {len(a)} tensors differ, first {bad[:5]}"


def test_training_state_survives_a_resize(gpu):
    """Weights, momentum (weight_updates after the update = momentum * gradient history), rolling statistics and seen:
    three steps with a resize to 320x352 and back before the second one == three steps without (bit for bit; the step is
    reproducible, tests/test_repeat_gpu.py)."""
    _same(_train_run(False), _train_run(True), "resize there and back")


def test_training_step_after_resize_equals_a_network_loaded_at_that_size(gpu):
    from darknet_b200 import lib as L, network as N
    B, classes = 4, 80
    w, h = 352, 320
    text0 = cfgs.CONFIGS["yolov4-tiny"](batch=B, subdivisions=1, w=416, h=416)
    weights = cfgmod.synth_weights(text0)
    flags = L.NET_DEFAULT | L.NET_TRAIN
    a = N.LoadSynthetic(text0, weights, batch=B, flags=flags)
    N.ResizeNetwork(a, w, h)
    b = N.LoadSynthetic(cfgs.CONFIGS["yolov4-tiny"](batch=B, subdivisions=1, w=w, h=h), weights, batch=B, flags=flags)
    outs = []
    for net in (a, b):
        o = {}
        for s in range(2):
            x = cfgmod.synth_images(B, 3, h, w, seed=15 + s)
            truth = cfgmod.synth_truth(B, 90, classes=classes, seed=16 + s)
            o[f"cost{s}"] = np.float32(N.TrainNetworkDatum(net, x, truth))
            assert np.isfinite(o[f"cost{s}"])
            N.UpdateNetwork(net)
            _snapshot(net, N, o, f"s{s}")
        outs.append(o)
        net.close()
    _same(outs[0], outs[1], "resized vs loaded at 352x320")


def test_shim_ResizeNetwork_keeps_the_reference_struct_in_step(gpu):
    """libyolo_core_b200.so: ResizeNetwork(Network*, w, h) with the reference's signature; the mirrored `Network` /
    `layer` fields follow, and NetworkPredict + GetNetworkBoxes on the resized struct == a network loaded at that size."""
    from darknet_b200 import network as N, shim
    from tests.netutil import materialise
    if not shim.have():
        pytest.skip("libyolo_core_b200.so not built")
    _, cpath, wpath = materialise("yolov4-tiny", head_bias=-3.0, head_gain=40.0)
    _, cpath2, _ = materialise("yolov4-tiny", head_bias=-3.0, head_gain=40.0, w=320, h=352)
    net = shim.ShimNetwork(cpath, wpath)
    net.resize(320, 352)
    x = cfgmod.synth_images(1, 3, 352, 320, seed=3)
    net.predict(x)
    rows, probs = net.boxes(0.25, 0.45, 80)
    own = N.LoadNetwork(cpath2, wpath, batch=1)
    N.NetworkPredict(own, x)
    d = N.GetNetworkBoxes(own, 0.25, 0)
    N.NmsSort(d, 0.45)
    assert rows.shape[0] > 0 and np.array_equal(rows, d.boxes()) and np.array_equal(probs, d.probs())
    own.close()
    net.close()
