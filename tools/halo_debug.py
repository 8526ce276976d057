"""Debug the halo conv mode: error per single active filter tap."""
import os, sys
os.environ["YC_CONV_HALO"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from darknet_b200 import ops
from oracle.pyoracle import O
gpu = torch.device("cuda:0")
def bf(a): return torch.from_numpy(a).to(torch.bfloat16).float().numpy()
B, C, H, W, N = 1, 64, 16, 8, 64
rng = np.random.default_rng(0)
x = bf(rng.standard_normal((B, C, H, W)).astype(np.float32))
wfull = bf((rng.standard_normal((N, C, 3, 3)) * 0.1).astype(np.float32))
b = np.zeros(N, np.float32)
xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
for tap in [1, 2, 4, -1]:
    w = np.zeros_like(wfull)
    if tap < 0: w[:] = wfull
    else: w[:, :, tap // 3, tap % 3] = wfull[:, :, tap // 3, tap % 3]
    ref = O.conv_forward(x, w, b, 3, 1, 1, "linear")
    conv = ops.Conv(xd, (B, H, W, C), torch.from_numpy(w), torch.from_numpy(b), 3, 1, 1, "linear")
    y = conv(); torch.cuda.synchronize()
    got = ops.to_nchw(y, N).cpu().numpy()
    err = np.abs(got - ref).max() / (np.abs(ref).max() + 1e-12)
    # where is it wrong? per output row / col error
    e_hw = np.abs(got - ref).max(axis=(0, 1))
    print("tap", tap, f"err {err:.3e}", "bad rows", np.nonzero(e_hw.max(axis=1) > 1e-2 * np.abs(ref).max())[0].tolist()[:20],
          "bad cols", np.nonzero(e_hw.max(axis=0) > 1e-2 * np.abs(ref).max())[0].tolist(), flush=True)
