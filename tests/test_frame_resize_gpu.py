"""cv::resize of ProcImage (yolo_core.cpp:108-109) on the device: yc_resize_u8 / yc_resize_u8_to_nchw against the oracle
(oracle/cv_resize.py, itself pinned bit for bit against cv2.resize by tests/golden/cv_resize.npz): integer work, so the bar
is bit-exact; and the frame entry point yc_network_predict_u8_resized == yc_network_predict_u8 on oracle-resized frames."""
import numpy as np
import pytest
import torch

from darknet_b200 import cfg as cfgmod
from oracle import cv_resize as R
from tests.netutil import materialise

pytestmark = pytest.mark.gpu

CASES = [((17, 5), (33, 64), 3), ((2, 2), (7, 9), 3), ((1, 1), (4, 4), 1), ((40, 37), (20, 37), 3), ((64, 64), (32, 32), 3),
         ((64, 64), (32, 31), 4), ((5, 70), (3, 10), 3), ((1, 9), (5, 3), 1), ((9, 1), (3, 5), 3), ((61, 61), (61, 61), 3),
         ((480, 640), (608, 608), 3), ((1080, 1920), (608, 608), 3), ((1216, 1216), (608, 608), 3), ((375, 500), (416, 416), 3)]


@pytest.mark.parametrize("ssize,dsize,c", CASES)
def test_resize_u8_bit_exact(gpu, ssize, dsize, c):
    from darknet_b200 import lib as L
    (sh, sw), (dh, dw) = ssize, dsize
    n = 3
    rng = np.random.default_rng(sh * 1000 + sw)
    src = rng.integers(0, 256, (n, sh, sw, c), dtype=np.uint8)
    want = np.stack([R.resize_u8(src[i], dh, dw).reshape(dh, dw, c) for i in range(n)])
    sd = torch.from_numpy(src).to(gpu)
    dd = torch.full((n, dh, dw, c), 77, dtype=torch.uint8, device=gpu)
    L.check(L.lib().yc_resize_u8(sd.data_ptr(), n, c, sh, sw, dd.data_ptr(), dh, dw, L.stream_ptr()), "yc_resize_u8")
    torch.cuda.synchronize()
    got = dd.cpu().numpy()
    assert np.array_equal(got, want), f"{int((got != want).sum())} of {got.size} bytes differ (max {np.abs(got.astype(int) - want).max()})"
    # fused with cvtColor(RGB2BGR) + Mat2Image: fp32 NCHW, value / 255.0f, channel c-1-k
    for swap in (0, 1):
        fd = torch.zeros((n, c, dh, dw), dtype=torch.float32, device=gpu)
        L.check(L.lib().yc_resize_u8_to_nchw(sd.data_ptr(), n, c, sh, sw, swap, fd.data_ptr(), dh, dw, L.stream_ptr()), "yc_resize_u8_to_nchw")
        torch.cuda.synchronize()
        w = want[..., ::-1] if swap else want
        wf = (w.astype(np.float32) / np.float32(255.0)).transpose(0, 3, 1, 2)
        assert np.array_equal(fd.cpu().numpy(), wf)


def test_predict_from_frames_equals_predict_on_resized(gpu):
    from darknet_b200 import network as N
    text, cpath, wpath = materialise("micro")
    B = 3
    net = N.LoadNetwork(cpath, wpath, batch=B)
    rng = np.random.default_rng(5)
    for (sh, sw) in ((48, 80), (128, 128), (64, 64), (200, 90)):       # up, exact 2 x 2 (area), same size, mixed
        frames = rng.integers(0, 256, (B, sh, sw, 3), dtype=np.uint8)
        N.NetworkPredictFrames(net, frames, swap_rb=True)
        got = {i: np.array(net.layer_output(i), copy=True) for i in net.yolo_layers() + [0]}
        resized = np.stack([R.resize_u8(frames[i], 64, 64) for i in range(B)])
        N.NetworkPredictU8(net, resized, swap_rb=True)
        for i, g in got.items():
            assert np.array_equal(g, net.layer_output(i)), f"frames {sh}x{sw}: layer {i} differs"
    net.close()
