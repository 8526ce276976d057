"""TEST / BENCH INFRASTRUCTURE — times the UNMODIFIED reference CPU path (oracle/_ref/libyolo_ref.so).

Run as a subprocess by bench.py (cpu_baseline leg and --impl reference) so that OMP_NUM_THREADS can be
fixed before the reference's OpenMP runtime starts.  Prints one JSON line:
  {"images_per_s": V, "s_per_image": T, "images": n, "threads": C, "config": name, "w":..,"h":..}
The sample is bounded: one warm-up NetworkPredict, then as many timed calls as fit `--budget` seconds
(at least `--min-images`).  Inference is the reference's native batch-1 path (parser.cpp:1114-1115).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="yolov4")
    ap.add_argument("--width", type=int, default=608)
    ap.add_argument("--height", type=int, default=608)
    ap.add_argument("--budget", type=float, default=20.0)
    ap.add_argument("--min-images", type=int, default=1)
    ap.add_argument("--max-images", type=int, default=64)
    ap.add_argument("--train", type=int, default=0, help="time TrainNetworkDatum+UpdateNetwork at this batch instead")
    a = ap.parse_args()
    import numpy as np
    from darknet_b200 import cfg as cfgmod, cfgs
    from oracle import pyoracle
    if not pyoracle.have_ref():
        print(json.dumps({"error": "oracle/_ref/libyolo_ref.so missing"}))
        return 0
    if a.train > 0:
        return train(a)
    text = cfgs.CONFIGS[a.config](w=a.width, h=a.height)
    d = tempfile.mkdtemp(prefix="yc_ref_")
    cpath, wpath = os.path.join(d, "n.cfg"), os.path.join(d, "n.weights")
    open(cpath, "w").write(text)
    cfgmod.write_weights(wpath, text, cfgmod.synth_weights(text))
    net = pyoracle.RefNet(cpath, wpath, train=False, fuse=True)
    x = cfgmod.synth_images(1, 3, a.height, a.width)
    t0 = time.perf_counter()
    net.time_predict(x, 1)  # warm-up
    warm = time.perf_counter() - t0
    t1 = net.time_predict(x, 1)
    n = int(max(a.min_images, min(a.max_images, a.budget / max(t1, 1e-6))))
    tt = net.time_predict(x, n) if n > 1 else t1
    per = tt / max(n, 1)
    print(json.dumps({"images_per_s": 1.0 / per, "s_per_image": per, "images": n, "warmup_s": warm,
                      "threads": int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1)), "config": a.config,
                      "w": a.width, "h": a.height}))
    return 0


def train(a):
    """Training step of the reference CPU path (ForwardNetwork(train) + BackwardNetwork + UpdateNetwork) at batch
    a.train; its backward uses the scalar gemm_nt / gemm_tn (gemm.c:2914-2947)."""
    from darknet_b200 import cfg as cfgmod, cfgs
    from oracle import pyoracle
    B = a.train
    text = cfgs.CONFIGS[a.config](w=a.width, h=a.height, batch=B, subdivisions=1)
    d = tempfile.mkdtemp(prefix="yc_ref_")
    cpath, wpath = os.path.join(d, "n.cfg"), os.path.join(d, "n.weights")
    open(cpath, "w").write(text)
    cfgmod.write_weights(wpath, text, cfgmod.synth_weights(text))
    net = pyoracle.RefNet(cpath, wpath, train=True)
    x = cfgmod.synth_images(B, 3, a.height, a.width)
    classes = 3 if a.config == "micro" else 80
    truth = cfgmod.synth_truth(B, 90, classes=classes)
    t0 = time.perf_counter()
    net.forward_backward(x, truth)
    net.L.ref_update(net.h)
    t1 = time.perf_counter() - t0
    n = int(max(1, min(a.max_images, a.budget / max(t1, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(n):
        net.forward_backward(x, truth)
        net.L.ref_update(net.h)
    per = (time.perf_counter() - t0) / n
    print(json.dumps({"images_per_s": B / per, "s_per_image": per / B, "images": n * B, "steps": n, "batch": B, "warmup_s": t1,
                      "threads": int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1)), "config": a.config,
                      "w": a.width, "h": a.height, "mode": "train"}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
