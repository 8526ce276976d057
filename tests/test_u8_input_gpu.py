"""u8 input path: Mat2Image (visualize.cpp:26-54) + NetworkPredict with the /255 and the HWC->NCHW change
on the device.  Mat2Image is restated here in numpy; NetworkPredictU8 must give bit-identical head outputs to
NetworkPredict on that restatement, with and without the RGB<->BGR swap of yolo_core.cpp:110."""
import numpy as np
import pytest

from tests import netutil

pytestmark = pytest.mark.gpu


def mat2image(img_u8):
    """image->data[k*w*h + y*w + x] = data[y*step + x*c + k] / 255.0f, per image of the batch"""
    return np.ascontiguousarray((img_u8.astype(np.float32) / np.float32(255.0)).transpose(0, 3, 1, 2))


@pytest.mark.parametrize("swap", [False, True])
def test_predict_u8_matches_mat2image(gpu, swap):
    from darknet_b200 import lib as L, network as N
    B = 3
    text, cpath, wpath = netutil.materialise("micro", batch=1)
    net = N.LoadNetwork(cpath, wpath, batch=B, flags=L.NET_DEFAULT)
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, size=(B, net.hh, net.w, 3), dtype=np.uint8)
    src = img[..., ::-1] if swap else img
    want_last = np.array(N.NetworkPredict(net, mat2image(src)), copy=True)
    want = [np.array(net.layer_output(i), copy=True) for i in range(net.n) if net.layer_type(i) == "yolo"]
    got_last = np.array(N.NetworkPredictU8(net, img, swap_rb=swap), copy=True)
    got = [np.array(net.layer_output(i), copy=True) for i in range(net.n) if net.layer_type(i) == "yolo"]
    assert np.array_equal(want_last, got_last)
    for a, b in zip(want, got):
        assert np.array_equal(a, b)
    # the pinned staging buffer is usable as a zero-copy producer target
    buf = N.InputBufferU8(net)
    buf[...] = img
    again = np.array(N.NetworkPredictU8(net, buf, swap_rb=swap), copy=True)
    assert np.array_equal(again, got_last)
    net.close()


def test_prefetch_gives_identical_results(gpu):
    """yc_network_prefetch: a stream of batches fed through the double-buffered input gives the same outputs as plain
    NetworkPredict calls, in order, including when a prefetched batch differs from the one before."""
    from darknet_b200 import cfg as cfgmod, lib as L, network as N
    B = 2
    text, cpath, wpath = netutil.materialise("micro", batch=1)
    net = N.LoadNetwork(cpath, wpath, batch=B, flags=L.NET_DEFAULT)
    batches = [cfgmod.synth_images(B, 3, net.hh, net.w, seed=100 + k) for k in range(4)]
    want = [np.array(N.NetworkPredict(net, b), copy=True) for b in batches]
    slots = [N.InputBufferSlot(net, 0), N.InputBufferSlot(net, 1)]
    slots[0][...] = batches[0]
    for k in range(4):
        if k + 1 < 4:                                  # announce the next batch, then compute the current one
            slots[(k + 1) % 2][...] = batches[k + 1]
            N.Prefetch(net, slots[(k + 1) % 2])
        got = np.array(N.NetworkPredict(net, slots[k % 2]), copy=True)
        assert np.array_equal(got, want[k]), f"batch {k}"
    # a prefetched batch that is never asked for must not leak into a later call
    N.Prefetch(net, slots[1])
    got = np.array(N.NetworkPredict(net, batches[0]), copy=True)
    assert np.array_equal(got, want[0])
    got = np.array(N.NetworkPredict(net, batches[2]), copy=True)
    assert np.array_equal(got, want[2])
    net.close()
