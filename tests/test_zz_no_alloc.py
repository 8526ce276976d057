"""yc_bias_act_bwd_noalloc (static device scratch, allocates nothing) against yc_bias_act_bwd (caller workspace).

The entry point has no counterpart in the reference (see
include/yc_b200.h).  Both entry points launch the same kernels on the same grid, so the comparison is exact: dz bit for
bit, bias_updates bit for bit (the sums are formed in a fixed order).  Two calls in a row on every case: the static
array keeps the previous call's contents and the second call must not see them.  Collected last on purpose
(a failure here hides nothing after it).
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(gpu, B, H, W, Cn, act, seed, ld_extra=0):
    from darknet_b200 import lib as L
    lib = L.lib()
    g = torch.Generator(device="cpu").manual_seed(seed)
    ld = (Cn + 7) // 8 * 8 + ld_extra
    dy = torch.zeros(B, H, W, ld, dtype=torch.bfloat16)
    y = torch.zeros(B, H, W, ld, dtype=torch.bfloat16)
    dy[..., :Cn] = (torch.randn(B, H, W, Cn, generator=g) * 0.1).to(torch.bfloat16)
    y[..., :Cn] = torch.randn(B, H, W, Cn, generator=g).to(torch.bfloat16)
    bu0 = torch.randn(Cn, generator=g)
    dy, y = dy.to(gpu), y.to(gpu)
    st = L.stream_ptr()
    dyv, yv = L.view(dy, B, H, W, Cn, ld), L.view(y, B, H, W, Cn, ld)

    # existing entry point: caller-allocated, zero-filled workspace
    ws = torch.zeros(lib.yc_bn_bwd_workspace_floats(Cn), dtype=torch.float32, device=gpu)
    dz_a = torch.full((B, H, W, ld), 7.0, dtype=torch.bfloat16, device=gpu)
    bu_a = bu0.clone().to(gpu)
    # new entry point: buffers created BEFORE the calls; nothing is allocated between or inside them
    dz_b = torch.full((B, H, W, ld), 7.0, dtype=torch.bfloat16, device=gpu)
    bu_b = bu0.clone().to(gpu)
    torch.cuda.synchronize()
    for call in range(2):  # two calls in a row: bias_updates accumulates, dz is rewritten
        L.check(lib.yc_bias_act_bwd(dyv, yv, L.ACT[act], bu_a.data_ptr(), ws.data_ptr(), L.view(dz_a, B, H, W, Cn, ld), st),
                "yc_bias_act_bwd")
        L.check(lib.yc_bias_act_bwd_noalloc(dyv, yv, L.ACT[act], bu_b.data_ptr(), L.view(dz_b, B, H, W, Cn, ld), st),
                "yc_bias_act_bwd_noalloc")
        torch.cuda.synchronize()
        assert torch.equal(dz_a.view(torch.int16), dz_b.view(torch.int16)), f"dz differs on call {call}"
        assert torch.equal(bu_a.view(torch.int32), bu_b.view(torch.int32)), \
            f"bias_updates differ on call {call}: max {float((bu_a - bu_b).abs().max()):.3e}"
    # and the sums are the right ones (float64 on the host, tolerance of the fp32 fixed-order sum)
    d, u = dy[..., :Cn].double().cpu(), y[..., :Cn].double().cpu()
    slope = {"leaky": torch.where(u > 0, 1.0, 0.1), "linear": torch.ones_like(u), "relu": (u > 0).double(),
             "logistic": u * (1 - u)}[act]
    want = bu0.double() + 2 * (d * slope).sum(dim=(0, 1, 2))
    assert float((bu_b.double().cpu() - want).abs().max()) <= 1e-4 * float(want.abs().max() + 1.0)


def test_noalloc_matches_workspace_entry_point(gpu):
    _run(gpu, 4, 26, 26, 255, "linear", 1)       # a YOLO head (no batchnorm, linear): what the training step runs
    _run(gpu, 2, 19, 19, 24, "leaky", 2)         # fewer channels than the call before: stale partials must not leak
    _run(gpu, 3, 13, 17, 1024, "leaky", 3)       # the stated maximum
    _run(gpu, 2, 8, 8, 40, "logistic", 4, ld_extra=8)  # pixel pitch wider than the channels


def test_noalloc_rejects_more_than_the_maximum(gpu):
    from darknet_b200 import lib as L
    lib = L.lib()
    cmax = lib.yc_bias_act_bwd_noalloc_max_channels()
    assert cmax == 1024
    Cn = cmax + 8
    dy = torch.zeros(1, 4, 4, Cn, dtype=torch.bfloat16, device=gpu)
    dz = torch.zeros_like(dy)
    bu = torch.zeros(Cn, device=gpu)
    rc = lib.yc_bias_act_bwd_noalloc(L.view(dy, 1, 4, 4, Cn), L.view(dy, 1, 4, 4, Cn), L.ACT["linear"], bu.data_ptr(),
                                     L.view(dz, 1, 4, 4, Cn), L.stream_ptr())
    assert rc == -2  # YC_ERR_ARG
    msg = lib.yc_last_error().decode()
    assert "exceed" in msg and "1024" in msg, msg
    torch.cuda.synchronize()
    assert float(bu.abs().max()) == 0.0
