"""Halo mode of the conv kernel (a_mode 2: one staged input patch per tile, nine shifted UMMA descriptors) against the
CPU oracle, forced on with YC_CONV_HALO=1; same tolerances as tests/test_conv_gpu.py.  Shapes exercise full tiles,
right/bottom clipping (H, W not multiples of 16 / 8), several 64-channel chunks, several N tiles, the fused residual and
fp32 outputs."""
import pytest

from tests.test_conv_gpu import _run

pytestmark = pytest.mark.gpu


@pytest.fixture
def halo(monkeypatch):
    monkeypatch.setenv("YC_CONV_HALO", "1")


@pytest.mark.parametrize("C,N,H,W,act", [(64, 64, 16, 8, "linear"), (64, 64, 32, 32, "leaky"), (128, 128, 19, 19, "mish"),
                                         (64, 96, 7, 9, "mish"), (256, 256, 13, 13, "leaky"), (128, 64, 38, 38, "mish"),
                                         (64, 32, 20, 76, "leaky"), (32, 64, 32, 24, "mish"), (32, 32, 19, 19, "leaky"),
                                         (96, 64, 16, 16, "leaky")])
def test_conv3x3_halo(gpu, halo, C, N, H, W, act):
    from darknet_b200 import ops
    _run(gpu, 2, C, H, W, N, 3, 1, act)


def test_conv3x3_halo_is_selected(gpu, halo):
    import numpy as np
    import torch
    from darknet_b200 import ops
    x = ops.to_nhwc(torch.zeros((1, 64, 16, 16), device=gpu))
    conv = ops.Conv(x, (1, 16, 16, 64), torch.zeros((64, 64, 3, 3)), torch.zeros(64), 3, 1, 1, "linear")
    assert conv.info()["a_mode"] == 2, conv.info()


def test_conv3x3_halo_residual_and_fp32(gpu, halo):
    _run(gpu, 2, 128, 19, 19, 128, 3, 1, "leaky", residual=True)
    _run(gpu, 2, 64, 24, 24, 48, 3, 1, "linear", out_f32=True)
