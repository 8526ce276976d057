"""Time the batchnorm / activation passes (yc_bn_act_fwd, yc_bn_bwd) alone on the layer shapes of yolov4-tiny 416 x 64.
usage: python tools/bn_bench.py [--act leaky]   -> per shape: us and GB/s of algorithmic bytes (fwd 2 tensors, bwd 5)"""
import argparse, ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from darknet_b200 import lib as L

ap = argparse.ArgumentParser()
ap.add_argument("--act", default="leaky")
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
lib = L.lib()
act = L.ACT[a.act] if hasattr(L, "ACT") else {"linear": 0, "leaky": 1, "mish": 2}[a.act]
shapes = [(208, 32), (104, 64), (104, 32), (52, 128), (52, 64), (26, 256), (26, 128), (13, 512), (13, 256)]
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = L.stream_ptr()
for hw, c in shapes:
    n = a.batch
    cd = lib.yc_conv_dgrad_channels(c)
    z = torch.randn(n, hw, hw, c, device=dev).to(torch.bfloat16)
    y = torch.empty_like(z)
    dy = torch.randn(n, hw, hw, c, device=dev).to(torch.bfloat16)
    dz = torch.empty(n, hw, hw, cd, device=dev, dtype=torch.bfloat16)
    mean = torch.zeros(c, device=dev); var = torch.ones(c, device=dev)
    gamma = torch.ones(c, device=dev); beta = torch.zeros(c, device=dev)
    su = torch.zeros(c, device=dev); ws = torch.zeros(lib.yc_bn_bwd_workspace_floats(c), device=dev)
    zv, yv, dyv, dzv = (L.view(t, n, hw, hw, t.shape[-1] if t is dz else c, t.shape[-1]) for t in (z, y, dy, dz))
    def fwd():
        L.check(lib.yc_bn_act_fwd(zv, mean.data_ptr(), var.data_ptr(), gamma.data_ptr(), beta.data_ptr(), act, yv, st), "fwd")
    def bwd():
        L.check(lib.yc_bn_bwd(dyv, zv, mean.data_ptr(), var.data_ptr(), gamma.data_ptr(), beta.data_ptr(), act,
                              su.data_ptr(), None, ws.data_ptr(), dzv, st), "bwd")
    out = []
    for fn, tensors in ((fwd, 2.0), (bwd, 4.0 + cd / c)):
        fn(); torch.cuda.synchronize()
        ts = []
        for _ in range(a.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        us = sorted(ts)[len(ts) // 2]
        out.append((us, tensors * z.numel() * 2 / us * 1e-3))
    print(f"{hw:4d}^2 x {c:4d} (dz pitch {cd:4d})  fwd {out[0][0]:7.1f} us {out[0][1]:7.0f} GB/s   bwd {out[1][0]:7.1f} us {out[1][1]:7.0f} GB/s")
