"""The gates of the reference's training loops, as cfg options of a layer:
  stopbackward=1            BackwardNetworkGpu stops BEFORE that layer (network_kernels.cu:140-143): it and every earlier
                            layer receive no gradient;
  dont_update=k             UpdateNetworkGpu skips the layer while dont_update >= iteration (network_kernels.cu:236-240);
  train_only_bn=1           marks that layer and every EARLIER one (SetTrainOnlyBn, parser.cpp:1062-1074); the update skips
                            them (network_kernels.cu:234-235).
The step is bit-reproducible (tests/test_repeat_gpu.py), so "unchanged" and "same as the ungated run" are exact."""
import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod, cfgs

pytestmark = pytest.mark.gpu

B, HW, CLASSES = 4, 64, 3


def _with_option(text, layer_index, line):
    """cfg text with `line` appended to the section of layer `layer_index` (sections after [net], in order)"""
    parts = text.split("\n[")
    # parts[0] = "[net]...", parts[1 + i] = section of layer i (without its leading '[')
    sec = parts[1 + layer_index].rstrip("\n")
    parts[1 + layer_index] = sec + "\n" + line + "\n"
    return "\n[".join(parts)


def _step(text, weights):
    from darknet_b200 import lib as L, network as N
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    convs = [i for i in range(net.n) if net.layer_info(i, 0) == 0]
    before = {i: N.layer_param(net, i, "weights") for i in convs}
    x = cfgmod.synth_images(B, 3, HW, HW, seed=21)
    truth = cfgmod.synth_truth(B, 90, classes=CLASSES, seed=22)
    cost = N.TrainNetworkDatum(net, x, truth)
    grads = {i: N.layer_param(net, i, "weight_updates") for i in convs}
    N.UpdateNetwork(net)
    after = {i: N.layer_param(net, i, "weights") for i in convs}
    net.close()
    return convs, cost, before, grads, after


def test_stopbackward_dont_update_train_only_bn(gpu):
    text = cfgs.micro(batch=B, subdivisions=1, classes=CLASSES)
    weights = cfgmod.synth_weights(text)
    convs, cost0, before0, grads0, after0 = _step(text, weights)
    assert all(np.abs(grads0[i]).max() > 0 for i in convs), "the ungated step reaches every conv layer"
    assert all(not np.array_equal(before0[i], after0[i]) for i in convs), "the ungated update moves every conv layer"

    # stopbackward=1 on layer 9: layers 0..9 get no gradient, later layers exactly the ungated one
    _, cost, _, grads, after = _step(_with_option(text, 9, "stopbackward=1"), weights)
    assert cost == cost0
    for i in convs:
        if i <= 9:
            assert np.abs(grads[i]).max() == 0.0, f"layer {i} received a gradient behind stopbackward"
        else:
            assert np.array_equal(grads[i], grads0[i]), f"layer {i}: gradient differs from the ungated step"

    # dont_update far in the future on layer 14: only that layer keeps its weights
    _, _, before, _, after = _step(_with_option(text, 14, "dont_update=100000"), weights)
    for i in convs:
        if i == 14:
            assert np.array_equal(before[i], after[i]), "dont_update layer was updated"
        else:
            assert np.array_equal(after[i], after0[i]), f"layer {i}: update differs from the ungated step"

    # train_only_bn on layer 6: layers 0..6 are skipped by the update, later ones as ungated
    _, _, before, _, after = _step(_with_option(text, 6, "train_only_bn=1"), weights)
    for i in convs:
        if i <= 6:
            assert np.array_equal(before[i], after[i]), f"layer {i} (train_only_bn range) was updated"
        else:
            assert np.array_equal(after[i], after0[i]), f"layer {i}: update differs from the ungated step"
