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


@pytest.mark.parametrize("stream_kind", ["null", "default", "own"])
def test_steps_enqueued_back_to_back_equal_synchronised_steps(gpu, stream_kind):
    """yc_network_forward_backward + yc_network_update enqueued with no host synchronisation in between (how a training
    loop and bench.py run them) give the same bits as the same calls with a synchronisation after each — on a NULL stream
    (= the network's own stream in BOTH entry points), on torch's default stream (handle 0, i.e. NULL again) and on a
    stream of the caller's.  Round 2 regression: NULL used to mean the legacy default stream in forward_backward and the
    network's stream in update, so the update raced with the backward pass before it."""
    import torch
    from darknet_b200 import lib as L, network as N
    B, hw = 8, 416
    text = cfgs.CONFIGS["yolov4-tiny"](batch=B, subdivisions=1, w=hw, h=hw)
    weights = cfgmod.synth_weights(text)
    xd = torch.from_numpy(cfgmod.synth_images(B, 3, hw, hw, seed=31)).to(gpu)
    td = torch.from_numpy(cfgmod.synth_truth(B, 90, classes=80, seed=32)).to(gpu)
    own = torch.cuda.Stream()
    torch.cuda.synchronize()
    sp = {"null": None, "default": C.c_void_p(torch.cuda.current_stream().cuda_stream), "own": C.c_void_p(own.cuda_stream)}[stream_kind]
    res = []
    for sync in (True, False):
        net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
        for _ in range(6):
            L.check(L.lib().yc_network_forward_backward(net.h, C.c_void_p(xd.data_ptr()), C.c_void_p(td.data_ptr()), sp), "fb")
            if sync:
                torch.cuda.synchronize()
            L.check(L.lib().yc_network_update(net.h, -1.0, -1.0, -1.0, sp), "update")
            if sync:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        res.append({i: N.layer_param(net, i, "weights") for i in range(net.n) if net.layer_info(i, 0) == 0})
        net.close()
    bad = [i for i in res[0] if not np.array_equal(res[0][i], res[1][i])]
    assert not bad, f"{stream_kind} stream: layers {bad[:6]} differ between synchronised and back-to-back steps"


def test_profile_pass_leaves_no_trace_in_the_training_state(gpu):
    """yc_network_train_profile times forward + backward + an update with lr = 0: afterwards the gradient accumulators
    (momentum history), the rolling batchnorm statistics, the weights and the counters are what they were (ADVICE round 1:
    the timing pass used to scale the accumulators by the momentum and move the rolling statistics)."""
    import torch
    from darknet_b200 import lib as L, network as N
    B, hw = 4, 64
    text = cfgs.micro(batch=B, subdivisions=1)
    net = N.LoadSynthetic(text, cfgmod.synth_weights(text), batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    x = cfgmod.synth_images(B, 3, hw, hw, seed=41)
    truth = cfgmod.synth_truth(B, 90, classes=3, seed=42)
    N.TrainNetworkDatum(net, x, truth)
    N.UpdateNetwork(net)

    def state():
        out = {"seen": np.int64(net.info(10))}
        for i in range(net.n):
            if net.layer_info(i, 0) == 0:
                for key in ("weights", "weight_updates", "bias_updates", "biases") + (("scale_updates", "rolling_mean", "rolling_variance", "scales") if net.layer_info(i, 9) else ()):
                    out[f"L{i}.{key}"] = N.layer_param(net, i, key)
        return out
    before = state()
    xd, td = torch.from_numpy(x).to(gpu), torch.from_numpy(truth).to(gpu)
    ms = np.zeros(7, np.float32)
    L.check(L.lib().yc_network_train_profile(net.h, C.c_void_p(xd.data_ptr()), C.c_void_p(td.data_ptr()), 2, ms.ctypes.data_as(C.c_void_p)),
            "yc_network_train_profile")
    assert ms[6] > 0
    _assert_same(before, state(), "training state across a profile pass")
    net.close()
