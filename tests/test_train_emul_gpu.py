"""The whole training step on the GPU against the bf16-EMULATING CPU walker (oracle/pytrain.py, emulate_bf16=True).

The walker is the reference's algorithm (its fp32 mode is pinned against the unmodified reference to 1e-5 by
tests/test_oracle_train_pin.py) with a round-to-bf16 inserted at exactly the points where the B200 path stores bf16.
Agreement here therefore says: the GPU step differs from the reference ONLY by those roundings — layer order, concat
placement, overwrite-vs-accumulate semantics (B6), accumulation into weight_updates (B7), no gradient for the BN shift
(B8), the epsilon placements (B1, B5), the unbiased variance (B3), the loss, the SGD update are all the reference's.

Tolerances: smooth activation (mish): every tensor <= 1.5e-2 rel-L2.  leaky: gradients <= 6e-2, because leaky' jumps
by 10x at 0 and the two sides sum in different orders (the walker's own fp32 mode differs from the reference by up to
4e-2 in the same quantity for the same reason — see the pin test)."""
import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod
from oracle import pytrain
from tests import netutil
from tests.netutil import rel_l2

pytestmark = pytest.mark.gpu


def _compare(net, st, tol_out, tol_grad, tag):
    from darknet_b200 import network as N
    worst = dict(out=0.0, delta=0.0, wu=0.0, su=0.0, bu=0.0, stat=0.0)
    for L in st.layers:
        i = L.index
        e = rel_l2(net.layer_output(i), st.out[i])
        worst["out"] = max(worst["out"], e)
        assert e < tol_out, f"{tag}: layer {i} ({L.kind}) output rel-L2 {e:.3e}"
        e = rel_l2(N.layer_delta(net, i), st.delta[i])
        worst["delta"] = max(worst["delta"], e)
        assert e < tol_grad, f"{tag}: layer {i} ({L.kind}) delta rel-L2 {e:.3e}"
        if i not in st.P:
            continue
        p = st.P[i]
        e = rel_l2(N.layer_param(net, i, "weight_updates"), p["weight_updates"].ravel())
        worst["wu"] = max(worst["wu"], e)
        assert e < tol_grad, f"{tag}: layer {i} weight_updates rel-L2 {e:.3e}"
        if L.bn:
            e = rel_l2(N.layer_param(net, i, "scale_updates"), p["scale_updates"])
            worst["su"] = max(worst["su"], e)
            assert e < tol_grad, f"{tag}: layer {i} scale_updates rel-L2 {e:.3e}"
            for key, v in (("mean", st.saved[i]["mean"]), ("variance", st.saved[i]["var"]),
                           ("rolling_mean", p["rolling_mean"]), ("rolling_variance", p["rolling_variance"])):
                e = rel_l2(N.layer_param(net, i, key), v)
                worst["stat"] = max(worst["stat"], e)
                assert e < 5e-3, f"{tag}: layer {i} {key} rel-L2 {e:.3e}"
            assert np.abs(N.layer_param(net, i, "bias_updates")).max() == 0.0
        else:
            e = rel_l2(N.layer_param(net, i, "bias_updates"), p["bias_updates"])
            worst["bu"] = max(worst["bu"], e)
            assert e < tol_grad, f"{tag}: layer {i} bias_updates rel-L2 {e:.3e}"
    return worst


def _run(name, B, hw, classes, tol_out, tol_grad, steps=2, **kw):
    from darknet_b200 import lib as L, network as N
    text, cpath, wpath = netutil.materialise(name, batch=B, subdivisions=1, **kw)
    W, _ = cfgmod.read_weights(wpath, text)
    st = pytrain.Step(text, W, emulate_bf16=True)
    net = N.LoadNetwork(cpath, wpath, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    for step in range(steps):
        x = cfgmod.synth_images(B, 3, hw, hw, seed=5 + 10 * step)
        truth = cfgmod.synth_truth(B, 90, classes=classes, seed=6 + 10 * step)
        cw = st.forward_backward(x, truth)
        cg = N.TrainNetworkDatum(net, x, truth)
        assert abs(cg - cw) <= 5e-3 * abs(cw), f"step {step}: cost {cg} vs walker {cw}"
        worst = _compare(net, st, tol_out, tol_grad, f"{name} step {step}")
        print(name, kw, "step", step, "cost", cg, cw, {k: f"{v:.2e}" for k, v in worst.items()})
        st.update()
        N.UpdateNetwork(net)
        for i, p in st.P.items():
            e = rel_l2(N.layer_param(net, i, "weights"), p["weights"].ravel())
            assert e < 2e-3, f"step {step}: layer {i} weights after update rel-L2 {e:.3e}"
    net.close()


def test_micro_mish_vs_bf16_walker(gpu):
    _run("micro", 4, 64, 3, 1.5e-2, 1.5e-2, act="mish")


def test_micro_leaky_vs_bf16_walker(gpu):
    _run("micro", 4, 64, 3, 1.5e-2, 6e-2, act="leaky")


def test_yolov4_tiny_vs_bf16_walker(gpu):
    """BASELINE config 4 at batch 2."""
    _run("yolov4-tiny", 2, 416, 80, 1.5e-2, 6e-2, steps=1, w=416, h=416)
