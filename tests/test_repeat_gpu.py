"""Reproducibility of the training step: no float atomics on the path (batchnorm statistics, batchnorm / bias backward
sums, split-K weight gradients and the loss sums are all formed in a fixed order), so two runs of the same step on fresh
instances give the same BITS — which is also what lets a data-parallel step be compared bit for bit with a local one.
Round 1 measured 15-30 % rel-L2 between two runs of yolov4-tiny's gradients (fp32 atomics + leaky' branch flips)."""
import ctypes as C

import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod, cfgs

pytestmark = pytest.mark.gpu


def _one_step(name, B, hw, classes, dp_world1=False, steps=2, **kw):
    from darknet_b200 import lib as L, network as N
    text = cfgs.CONFIGS[name](batch=B, subdivisions=1, **kw)
    weights = cfgmod.synth_weights(text)
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    if dp_world1:   # a one-rank NCCL world: the per-layer all-reduce path runs, on identity data
        uid = (C.c_ubyte * 128)()
        L.check(L.lib().yc_dp_unique_id(uid), "yc_dp_unique_id")
        L.check(L.lib().yc_network_dp_init(net.h, uid, 0, 1), "yc_network_dp_init")
        assert L.lib().yc_network_dp_info(net.h, 0) == 1 and L.lib().yc_network_dp_info(net.h, 2) > 0
    out = {}
    for s in range(steps):
        x = cfgmod.synth_images(B, 3, hw, hw, seed=5 + s)
        truth = cfgmod.synth_truth(B, 90, classes=classes, seed=6 + s)
        out[f"cost{s}"] = np.float32(N.TrainNetworkDatum(net, x, truth))
        for i in range(net.n):
            t = net.layer_info(i, 0)
            if t == 0:
                for key in ("weight_updates", "bias_updates") + (("scale_updates", "mean", "variance") if net.layer_info(i, 9) else ()):
                    out[f"s{s}.L{i}.{key}"] = N.layer_param(net, i, key)
            if t in (0, 4):
                out[f"s{s}.L{i}.delta"] = N.layer_delta(net, i)
        N.UpdateNetwork(net)
        for i in range(net.n):
            if net.layer_info(i, 0) == 0:
                out[f"s{s}.L{i}.weights"] = N.layer_param(net, i, "weights")
    net.close()
    return out


def _assert_same(a, b, what):
    assert a.keys() == b.keys()
    bad = [k for k in a if not np.array_equal(a[k], b[k])]
    assert not bad, f"{what}: {len(bad)} of {len(a)} tensors differ, first {bad[:5]}"


@pytest.mark.parametrize("name,B,hw,classes,kw", [("micro", 4, 64, 3, {}), ("micro", 4, 64, 3, {"gaussian": True}),
                                                  ("yolov4-tiny", 8, 416, 80, {"w": 416, "h": 416})])
def test_training_step_is_bit_reproducible(gpu, name, B, hw, classes, kw):
    a = _one_step(name, B, hw, classes, **kw)
    b = _one_step(name, B, hw, classes, **kw)
    _assert_same(a, b, f"{name}: two runs of the same two training steps")


def test_data_parallel_world1_equals_local_step(gpu):
    """yc_network_dp_init + the overlapped per-layer all-reduce in a one-rank world: the same bits as the local step
    (TrainNetworks / SyncNetworks replacement, network_kernels.cu:366-484; the multi-rank arithmetic — momentum carried
    as momentum / world, update by the global batch — is covered on the CPU by tests/test_dp_cpu.py and on 2+ GPUs by
    bench.py's training legs)."""
    a = _one_step("yolov4-tiny", 8, 416, 80, w=416, h=416)
    b = _one_step("yolov4-tiny", 8, 416, 80, dp_world1=True, w=416, h=416)
    _assert_same(a, b, "data-parallel world 1 vs local")
