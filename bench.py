#!/usr/bin/env python
"""bench.py — YOLOv4-608 batch-64 inference throughput on B200 (BASELINE.json metric), one JSON line.

  python bench.py --gpus N --steps K --warmup W            # our arm (libyc_b200.so through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU path (oracle/_ref)

A "step" = one forward pass of the hot path over one batch of 64 synthetic 608x608 images per GPU
(ForwardNetworkGpu: input layout conversion, 110 fused conv kernels, SPP maxpools, upsamples, YOLO decode).
  value : images/s, whole job, inputs already resident in HBM, timed with CUDA events on the launch stream,
          barrier + synchronize on both sides, max over ranks.
  e2e   : images/s through the public NetworkPredict-style calls with HOST buffers: every step copies one batch
          from pinned host memory (H2D; yc_network_prefetch starts the copy of the next batch while the current
          forward runs), runs the forward, extracts detections on the device and reads them back (D2H) + host NMS.
          e2e.serial = the same without prefetch (copy, then compute); e2e.u8_input = u8 HWC images (4x fewer bytes).
  roofline     : the conv kernel (tensor bound): algorithmic FLOPs (2*nweights*Ho*Wo per layer, the
                 reference's own bflops, convolutional_layer.cpp:714) / summed conv-kernel time measured with
                 CUDA events per launch; peak = MEASURED_PEAKS.json (sustained bf16, since the kernel is
                 timed inside a long step).
  cpu_baseline : the UNMODIFIED reference CPU path (oracle/_ref) on this box's host cores, bounded sample.
Multi-GPU: image-sharded, one process per GPU (torchrun), no collective on the data path -> weak scaling.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "images_per_sec"
UNIT = "images/s"
TRAIN_WL = "yolov4-tiny.cfg 416x416 training fwd+bwd+update batch 64, data-parallel NCCL grad allreduce"
TRAIN_WL_G = "yolov3 Gaussian-YOLO (gaussian_yolo_layer) 608x608 training fwd+bwd+update batch 64"
WORKLOADS = {
    "yolov4": ("yolov4", 608, "yolov4.cfg 608x608 batch 64 inference"),
    "yolov3": ("yolov3", 416, "yolov3.cfg 416x416 batch 64 inference"),
    "yolov4-tiny": ("yolov4-tiny", 416, "yolov4-tiny.cfg 416x416 inference"),
}


_REAL_STDOUT = None


def _quiet_stdout():
    """Route fd 1 to stderr for the duration of the run (NCCL prints 'NCCL version ...' on stdout when the first
    communicator comes up); emit() writes the one JSON line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, line)


def _ncu_traffic(n_conv):
    """DRAM bytes per conv launch from the committed ncu launch list of this same command (profiles/; ncu cannot run
    inside a timed bench).  None when the file is absent or does not hold a whole step."""
    import csv
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", "r[0-9][0-9]_launches_yolov4_608_b64.csv")))
    if not found:
        return None
    path = found[-1]  # the newest round's list
    try:
        with open(path, newline="") as f:
            lines = [l for l in f if l.startswith('"')]
        per = {}
        for r in csv.DictReader(lines):
            conv = "conv_fprop" in r["Kernel Name"] or "stem_conv" in r["Kernel Name"]  # (the stem is a conv launch of its own)
            if not conv or not r["Metric Name"].startswith("dram__bytes"):
                continue
            sc = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r["Metric Unit"], 1.0)
            per[int(r["ID"])] = per.get(int(r["ID"]), 0.0) + float(r["Metric Value"].replace(",", "")) * sc
        ids = sorted(per)
        if len(ids) < n_conv:
            return None
        last = ids[-n_conv:]
        return {"bytes_per_launch": sum(per[i] for i in last) / n_conv,
                "source": f"profiles/{os.path.basename(path)} (ncu launch list of bench.py --steps 1 --warmup 3, last step)"}
    except Exception:
        return None


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"burst": d.get("bf16_tflops", 1590.0), "sustained": d.get("bf16_tflops_sustained", 1400.0),
                "hbm": d.get("hbm_gbs", 6650.0), "source": "measured (MEASURED_PEAKS.json)"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cfgname, hw, wl = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    # each step = a bounded sample: `images_per_step` batch-1 NetworkPredict calls of the same workload
    per_step_budget = max(2.0, 150.0 / max(1, args.steps + args.warmup))
    env = dict(os.environ, OMP_NUM_THREADS=str(cores))
    vals = []
    info = {}
    for i in range(args.warmup + args.steps):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "time_ref.py"), "--config", cfgname, "--width", str(hw),
                            "--height", str(hw), "--budget", str(per_step_budget), "--max-images", "8"],
                           capture_output=True, text=True, env=env)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if not line:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref run failed: " + r.stderr[-200:]}))
            return 0
        info = json.loads(line[-1])
        if "error" in info:
            print(json.dumps({"impl": "reference", "unavailable": info["error"]}))
            return 0
        if i >= args.warmup:
            vals.append(info["images_per_s"])
    v = statistics.mean(vals)
    sample = f"{info['images']} batch-1 NetworkPredict calls per step after 1 warm-up (reference inference is batch 1, parser.cpp:1114-1115)"
    out = {"metric": METRIC, "value": v, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * 64 / v, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": wl, "net": cfgname, "width": hw, "height": hw, "batch": 64,
                      "note": "CPU reference timed per image; ms_per_step = 64 images at that rate"},
           "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
           "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))
    return 0


def _time_ref_train(cfgname, hw, budget, batch):
    cores = os.cpu_count() or 1
    env = dict(os.environ, OMP_NUM_THREADS=str(cores))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "time_ref.py"), "--config", cfgname, "--width", str(hw),
                        "--height", str(hw), "--budget", str(budget), "--max-images", "4", "--train", str(batch)],
                       capture_output=True, text=True, env=env, timeout=900)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    if not line:
        return None, cores, "oracle/_ref training run failed: " + r.stderr[-200:]
    info = json.loads(line[-1])
    if "error" in info:
        return None, cores, info["error"]
    return info, cores, (f"{info['steps']} step(s) of TrainNetworkDatum+UpdateNetwork at batch {info['batch']} "
                         f"({info['s_per_image']:.3f} s/image) after 1 warm-up step, OMP_NUM_THREADS={cores}; the reference's "
                         f"CPU backward uses its scalar gemm_nt/gemm_tn (gemm.c:2914-2947)")


def run_reference_train(args):
    """--impl reference --mode train: the reference's own CPU training step (forward(train) + backward + update)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    cfgname = args.train_net
    hw = 416 if cfgname == "yolov4-tiny" else 608
    vals, sample, cores = [], "", os.cpu_count() or 1
    per_step_budget = max(2.0, 150.0 / max(1, args.steps + args.warmup))
    for i in range(args.warmup + args.steps):
        info, cores, sample = _time_ref_train(cfgname, hw, per_step_budget, 8 if hw == 416 else 2)
        if info is None:
            print(json.dumps({"impl": "reference", "unavailable": sample}))
            return 0
        if i >= args.warmup:
            vals.append(info["images_per_s"])
    v = statistics.mean(vals)
    out = {"metric": "train_images_per_sec", "value": v, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * args.batch / v, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": TRAIN_WL if hw == 416 else TRAIN_WL_G, "net": cfgname, "width": hw, "height": hw, "batch": args.batch,
                      "note": "CPU reference timed at a reduced batch per step (bounded sample); ms_per_step = the named batch at that rate"},
           "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
           "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(out))
    return 0


def run_train(args):
    """--mode train: BASELINE configs[3].  A step = yc_network_forward_backward (forward with batch statistics, YOLO loss,
    backward with per-layer NCCL all-reduce overlapped when N > 1) + yc_network_update (SGD with momentum/decay) on one
    batch of 64 synthetic 416x416 images and 90 truth boxes per image per GPU."""
    import numpy as np
    import torch
    from darknet_b200 import cfg as cfgmod, cfgs, dp, lib as L, network as N

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU: the product has no CPU fallback"
    torch.cuda.set_device(local)
    L.check(L.lib().yc_set_device(local), "yc_set_device")
    dist = None
    if world > 1:
        import torch.distributed as dist  # noqa: F811
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfgname, B = args.train_net, args.batch
    if args.wgrad_single_tap:
        L.lib().yc_conv_wgrad_set_tap_groups(0)
    hw = 416 if cfgname == "yolov4-tiny" else 608
    text = cfgs.CONFIGS[cfgname](w=hw, h=hw, batch=B, subdivisions=1)
    weights = cfgmod.synth_weights(text)
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    if dist is not None:
        dp.init(net, dist)
    flops_fwd = net.flops_per_image()
    conv0 = [i for i in range(net.n) if net.layer_info(i, 0) == 0][0]
    f0 = 2.0 * net.layer_info(conv0, 1) * net.layer_info(conv0, 5) * net.layer_info(conv0, 6) ** 2 * net.layer_info(conv0, 2) * net.layer_info(conv0, 3)
    flops_step = (3.0 * flops_fwd - f0) * B      # fprop + wgrad for every conv, dgrad for all but the first
    x = cfgmod.synth_images(B, 3, hw, hw, seed=1234 + rank)
    truth = cfgmod.synth_truth(B, 90, classes=80, seed=4321 + rank)
    info_mib = L.lib().yc_network_train_info(net.h, 2)
    xd, td = torch.from_numpy(x).cuda(), torch.from_numpy(truth).cuda()
    # an explicit stream of our own for the whole step AND its timing events (torch's default stream is the legacy stream,
    # handle 0, which the library's training entry points read as "the network's own stream")
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    sp = C.c_void_p(stream.cuda_stream)

    def step():
        N.ForwardBackwardNetworkGpu(net, xd.data_ptr(), td.data_ptr(), sp)
        N.UpdateNetwork(net, stream=sp)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    L.lib().yc_launch_count(1)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = int(L.lib().yc_launch_count(0))
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * B * args.steps / (ms / 1e3)
    return _run_train_tail(args, locals())


def _run_train_tail(args, v):
    """e2e / roofline / cpu_baseline legs and the JSON line of --mode train (v = run_train's locals)."""
    import numpy as np
    import torch
    from darknet_b200 import lib as L, network as N
    net, dist, rank, world, B, hw = v["net"], v["dist"], v["rank"], v["world"], v["B"], v["hw"]
    x, truth, xd, td = v["x"], v["truth"], v["xd"], v["td"]
    e2e = roof = cpu = None
    if not args.quick_exit_after_timed:
        x_pin = N.InputBuffer(net)                   # pinned staging buffer (net->input_pinned_cpu): no pageable copy
        x_pin[...] = x
        def e2e_step():
            c = N.TrainNetworkDatum(net, x_pin, truth)   # pinned host fp32 NCHW + truth -> H2D -> forward/backward -> cost D2H
            N.UpdateNetwork(net)
            return c
        for _ in range(2):
            e2e_step()
        v["barrier"]()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            cost = e2e_step()
        torch.cuda.synchronize()
        t_e2e = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([t_e2e], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_e2e = float(t.item())
        e2e = {"value": world * B * args.steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(x.nbytes + truth.nbytes),
               "d2h_bytes_per_step": 4 * net.info(7), "ms_per_step": 1e3 * t_e2e / args.steps, "last_cost": cost,
               "api": "yc_network_train_datum(host fp32 NCHW images, host truth) + yc_network_update"}
        # every rank runs the profile pass (it contains the data-parallel all-reduces); rank 0 reports
        ms7 = np.zeros(7, np.float32)
        L.check(L.lib().yc_network_train_profile(net.h, C.c_void_p(xd.data_ptr()), C.c_void_p(td.data_ptr()), 3,
                                                 ms7.ctypes.data_as(C.c_void_p)), "yc_network_train_profile")
        if rank == 0:
            peaks = _peaks()
            conv_ms = float(ms7[0] + ms7[1] + ms7[2])
            ach = v["flops_step"] / (conv_ms / 1e3) / 1e12
            roof = {"bound": "tensor", "kernel": "conv_fprop_kernel (fprop + dgrad) + conv_wgrad_kernel", "achieved": ach,
                    "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": ach / peaks["sustained"], "traffic": None,
                    "peak_source": peaks["source"] + ", sustained bf16", "flops_per_step": v["flops_step"],
                    "ms_by_category": {"conv_fprop": float(ms7[0]), "conv_wgrad": float(ms7[1]), "conv_dgrad": float(ms7[2]),
                                       "bn_activation_passes": float(ms7[3]), "other_layers_and_loss": float(ms7[4]),
                                       "sgd_update": float(ms7[5]), "total_with_event_gaps": float(ms7[6])},
                    "how": "algorithmic FLOPs (fprop + wgrad of every conv, dgrad of all but the first) / summed duration of "
                           "those kernels, each operator bracketed by CUDA events (yc_network_train_profile, 3 passes)"}
            if not args.no_cpu_baseline and world == 1:
                info, cores, sample = _time_ref_train(v["cfgname"], hw, args.cpu_budget, 8 if hw == 416 else 2)
                cpu = {"value": info["images_per_s"] if info else None, "unit": UNIT, "cores": cores, "kind": "reference",
                       "sample": sample}
    if rank == 0:
        out = {"metric": "train_images_per_sec", "value": v["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
               "warmup": max(3, args.warmup), "ms_per_step": v["ms_per_step"], "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
               "config": {"workload": TRAIN_WL if v["cfgname"] == "yolov4-tiny" else TRAIN_WL_G, "net": v["cfgname"], "width": hw,
                          "height": hw, "batch_per_gpu": B,
                          "global_batch": B * world, "truth_boxes_per_image": 90,
                          "parallelism": (f"data-parallel x{world}: one NCCL all-reduce per conv-layer gradient bucket, overlapped "
                                          f"with backward ({L.lib().yc_network_dp_info(net.h, 2)} buckets, "
                                          f"{L.lib().yc_network_dp_info(net.h, 3)} MiB per step)") if world > 1 else "single GPU",
                          "optimizer": "SGD momentum .9 decay 5e-4 (cfg), fp32 master weights, bf16 operands",
                          "conv_gflop_per_image_fwd": v["flops_fwd"] / 1e9,
                          "l2_policy": "per-step working set (activations + deltas of 64 images, >2 GB) exceeds the 126 MB L2; no explicit flush"},
               "clocks": v["clocks"], "e2e": e2e, "gpu_launches": v["launches"], "roofline": roof, "cpu_baseline": cpu}
        emit(out)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0



def train_leg(rank, world, dist, cfgname, hw, B, steps, warmup=3):
    """Data-parallel training measured in the same run as the inference headline (BASELINE configs[3] / configs[4]): one
    process per GPU, B images per GPU per step (weak scaling), per-conv-layer gradient buckets all-reduced over NCCL on a
    separate stream while the rest of backward runs.  Reports the step (forward + backward + update), the same step with
    the exchange deferred (buckets stay local during backward: what backward costs without any communication), and
    their difference = the all-reduce time that backward could NOT hide."""
    import numpy as np
    import torch
    from darknet_b200 import cfg as cfgmod, cfgs, dp, lib as L, network as N
    lib = L.lib()
    text = cfgs.CONFIGS[cfgname](w=hw, h=hw, batch=B, subdivisions=1)
    weights = cfgmod.synth_weights(text)
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    del weights
    if dist is not None:
        dp.init(net, dist)
    x = torch.from_numpy(cfgmod.synth_images(B, 3, hw, hw, seed=1234 + rank)).cuda()
    t = torch.from_numpy(cfgmod.synth_truth(B, 90, classes=80, seed=4321 + rank)).cuda()
    stream = torch.cuda.Stream()   # explicit stream for the step and its events (see run_train)
    torch.cuda.synchronize()
    sp = C.c_void_p(stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n):
            fn()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1) / n
        if dist is not None:
            tt = torch.tensor([ms], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms

    def step():
        N.ForwardBackwardNetworkGpu(net, x.data_ptr(), t.data_ptr(), sp)
        N.UpdateNetwork(net, stream=sp)

    def fb():   # a backward pass without an update (timing only): declare the previous pass's reduced gradients discarded
        if dist is not None:
            dp.discard_step(net)
        N.ForwardBackwardNetworkGpu(net, x.data_ptr(), t.data_ptr(), sp)

    ms_step = timed(step, steps)
    out = {"net": cfgname, "width": hw, "height": hw, "batch_per_gpu": B, "global_batch": B * world, "n_gpus": world,
           "ms_per_step": ms_step, "images_per_sec": world * B / (ms_step / 1e3), "steps": steps, "scaling": "weak",
           "step": "yc_network_forward_backward + yc_network_update, device-resident synthetic batch, CUDA events, max over ranks"}
    if dist is not None:
        ms_fb = timed(fb, steps)                 # per-layer all-reduce overlapped with backward
        dp.set_overlap(net, False)
        dp.discard_step(net)
        def fb_local():                          # buckets stay local during backward; exchanged outside the compared region
            dp.discard_step(net)
            N.ForwardBackwardNetworkGpu(net, x.data_ptr(), t.data_ptr(), sp)
        for _ in range(warmup):
            fb_local()
            dp.reduce(net, sp)
        barrier()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2 * steps)]
        for k in range(steps):
            e[2 * k].record(stream)
            fb_local()
            e[2 * k + 1].record(stream)
            dp.reduce(net, sp)
        barrier()
        ms_local = sum(e[2 * k].elapsed_time(e[2 * k + 1]) for k in range(steps)) / steps
        tt = torch.tensor([ms_local], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_local = float(tt.item())
        dp.discard_step(net)
        dp.set_overlap(net, True)
        out.update({"buckets_per_step": lib.yc_network_dp_info(net.h, 2), "allreduce_mib_per_step": lib.yc_network_dp_info(net.h, 3),
                    "forward_backward_ms_overlapped": ms_fb, "forward_backward_ms_local_only": ms_local,
                    "exposed_allreduce_ms": ms_fb - ms_local,
                    "collective": "ncclAllReduce(sum, fp32) per conv-layer bucket [weight_updates | bias_updates | scale_updates] on a "
                                  "communication stream, issued as each layer's weight gradient is enqueued"})
    net.close()
    del net
    torch.cuda.empty_cache()
    return out


def cpu_baseline(cfgname, hw, budget):
    cores = os.cpu_count() or 1
    env = dict(os.environ, OMP_NUM_THREADS=str(cores))
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "time_ref.py"), "--config", cfgname, "--width", str(hw),
                            "--height", str(hw), "--budget", str(budget), "--max-images", "16"],
                           capture_output=True, text=True, env=env, timeout=600)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        info = json.loads(line[-1])
        if "error" in info:
            return {"value": None, "unit": UNIT, "cores": cores, "kind": "reference", "sample": info["error"]}
        return {"value": info["images_per_s"], "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"{info['images']} images, batch-1 NetworkPredict of the same net ({info['s_per_image']:.3f} s/image), "
                          f"OMP_NUM_THREADS={cores}, 1 warm-up"}
    except Exception as e:  # noqa: BLE001
        return {"value": None, "unit": UNIT, "cores": cores, "kind": "reference", "sample": f"failed: {e}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="infer: the headline (BASELINE configs[2]); train: configs[3], yolov4-tiny 416 fwd+bwd+update")
    ap.add_argument("--workload", default="yolov4", choices=sorted(WORKLOADS))
    ap.add_argument("--train-net", default="yolov4-tiny", choices=["yolov4-tiny", "gaussian-yolov3"],
                    help="--mode train: yolov4-tiny 416 (configs[3]) or Gaussian-YOLOv3 608 (configs[4])")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--layers", action="store_true", help="print the per-layer time table to stderr")
    ap.add_argument("--quick-exit-after-timed", action="store_true", help="(profiling) skip e2e / roofline / cpu legs")
    ap.add_argument("--net-flags", default="0", help="(A/B) extra yc_network_load flags in hex, OR'ed onto the default plan")
    ap.add_argument("--no-train-leg", action="store_true", help="skip the data-parallel training measurements of the default run")
    ap.add_argument("--wgrad-single-tap", action="store_true", help="(A/B, --mode train) plan wgrad with one tap per work unit")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_train(args) if args.mode == "train" else run_reference(args)
    _quiet_stdout()
    if args.mode == "train":
        return run_train(args)

    import numpy as np
    import torch
    from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU: the product has no CPU fallback"
    torch.cuda.set_device(local)
    L.check(L.lib().yc_set_device(local), "yc_set_device")
    dist = None
    if world > 1:
        import torch.distributed as dist  # noqa: F811
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    cfgname, hw, wl = WORKLOADS[args.workload]
    B = args.batch
    text = cfgs.CONFIGS[cfgname](w=hw, h=hw)
    weights = cfgmod.synth_weights(text)
    # sparse-detection regime like a trained detector: spread head logits, objectness bias -6
    cfgmod.shift_head_objectness(text, weights, -6.0, 40.0)
    flags = (L.NET_DEFAULT & ~L.NET_PULL_HEADS) | int(args.net_flags, 16)
    net = N.LoadSynthetic(text, weights, batch=B, flags=flags)
    del weights
    flops_img = net.flops_per_image()
    kernels_per_forward = net.kernels_per_forward()

    # synthetic images, different per rank (image sharding), resident on the device for the `value` leg
    x_host = N.InputBuffer(net)
    x_host[...] = cfgmod.synth_images(B, 3, hw, hw, seed=1234 + rank)
    x_dev = torch.from_numpy(np.ascontiguousarray(x_host)).cuda()
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)

    def step():
        L.check(L.lib().yc_network_forward(net.h, C.c_void_p(x_dev.data_ptr()), sp), "yc_network_forward")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    L.lib().yc_launch_count(1)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = int(L.lib().yc_launch_count(0))
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * B * args.steps / (ms / 1e3)

    # ---- e2e: host buffers, H2D + forward + on-device extraction + D2H + host NMS -------------------
    e2e = None
    roof = None
    cpu = None
    if not args.quick_exit_after_timed:
        def e2e_step():
            N.NetworkPredict(net, x_host)            # pinned host -> H2D -> forward (synchronous)
            return N.DetectBatch(net, 0.25, 0.45)    # device threshold/decode -> D2H rows -> host NmsSort
        for _ in range(2):
            rows, counts = e2e_step()
        barrier()
        t0 = time.perf_counter()
        d2h = 0
        for _ in range(args.steps):
            rows, counts = e2e_step()
            d2h += rows.nbytes + counts.nbytes
        torch.cuda.synchronize()
        t_e2e = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([t_e2e], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_e2e = float(t.item())
        e2e = {"value": world * B * args.steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(x_host.nbytes),
               "d2h_bytes_per_step": int(d2h // args.steps), "ms_per_step": 1e3 * t_e2e / args.steps,
               "detections_per_step": int(counts.sum()),
               "api": "yc_network_predict(pinned host fp32 NCHW) + yc_network_detect(thresh .25, nms .45)"}

        # the same step with the input double-buffered: yc_network_prefetch(next batch) right before
        # yc_network_predict(current batch), so the H2D of the next batch overlaps the forward of this one.  Every step
        # still moves one full batch host -> device inside the timed region.
        slots = [N.InputBufferSlot(net, 0), N.InputBufferSlot(net, 1)]
        slots[1][...] = slots[0]
        def e2e_pipe_step(k):
            N.Prefetch(net, slots[(k + 1) % 2])
            N.NetworkPredict(net, slots[k % 2])
            return N.DetectBatch(net, 0.25, 0.45)
        for k in range(2):
            e2e_pipe_step(k)
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            e2e_pipe_step(k)
        torch.cuda.synchronize()
        t_pipe = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([t_pipe], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_pipe = float(t.item())
        e2e["prefetch"] = {"value": world * B * args.steps / t_pipe, "unit": UNIT, "h2d_bytes_per_step": int(x_host.nbytes),
                           "ms_per_step": 1e3 * t_pipe / args.steps,
                           "api": "yc_network_prefetch(next pinned batch) + yc_network_predict(current) + yc_network_detect"}

        # same end-to-end step fed with u8 HWC images (ProcImage's Mat2Image + NetworkPredict): the H2D
        # copy carries bytes instead of floats
        u8 = N.InputBufferU8(net)
        u8[...] = np.clip(np.rint(np.asarray(x_host).transpose(0, 2, 3, 1) * 255.0), 0, 255).astype(np.uint8)
        def e2e_u8_step():
            N.NetworkPredictU8(net, u8)
            return N.DetectBatch(net, 0.25, 0.45)
        for _ in range(2):
            e2e_u8_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            rows8, counts8 = e2e_u8_step()
        torch.cuda.synchronize()
        t_u8 = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([t_u8], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_u8 = float(t.item())
        e2e["u8_input"] = {"value": world * B * args.steps / t_u8, "unit": UNIT, "h2d_bytes_per_step": int(u8.nbytes),
                           "ms_per_step": 1e3 * t_u8 / args.steps,
                           "api": "yc_network_predict_u8(pinned host u8 HWC) + yc_network_detect"}

        # the same double-buffered step in a DENSE detection regime (thousands of boxes per image stress the
        # extraction, the D2H of the rows and the host NMS): the threshold is set to the objectness of the batch's
        # 64000th strongest cell, i.e. about 1000 candidates per image (chosen outside the timed region)
        dense = None
        N.NetworkPredict(net, slots[0])
        probe, _ = N.DetectBatch(net, 0.02, 0.0)           # every cell with objectness > 0.02, no NMS
        obj_sorted = np.sort(probe[:, 5])[::-1]
        th = float(obj_sorted[min(len(obj_sorted) - 1, 1000 * B)]) if len(obj_sorted) else 0.02   # ~1000 candidates per image
        del probe
        def e2e_dense_step(k):
            N.Prefetch(net, slots[(k + 1) % 2])
            N.NetworkPredict(net, slots[k % 2])
            return N.DetectBatch(net, th, 0.45)
        for k in range(2):
            e2e_dense_step(k)
        barrier()
        t0 = time.perf_counter()
        d2h_d = 0
        n_dense = min(args.steps, 10)
        for k in range(n_dense):
            rows_d, counts_d = e2e_dense_step(k)
            d2h_d += rows_d.nbytes + counts_d.nbytes
        torch.cuda.synchronize()
        t_dense = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([t_dense], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_dense = float(t.item())
        dense = {"value": world * B * n_dense / t_dense, "unit": UNIT, "ms_per_step": 1e3 * t_dense / n_dense, "thresh": th,
                 "detections_per_step": int(counts_d.sum()), "d2h_bytes_per_step": int(d2h_d // n_dense),
                 "h2d_bytes_per_step": int(x_host.nbytes), "steps": n_dense,
                 "api": "yc_network_prefetch + yc_network_predict + yc_network_detect(dense thresh, nms .45)"}

        # headline e2e = the double-buffered stream (what a serving loop calls); the strictly serial call sequence and
        # the u8-input variant are reported beside it
        serial = {k: e2e[k] for k in ("value", "unit", "ms_per_step", "api")}
        e2e = {"value": e2e["prefetch"]["value"], "unit": UNIT, "h2d_bytes_per_step": e2e["h2d_bytes_per_step"],
               "d2h_bytes_per_step": e2e["d2h_bytes_per_step"], "ms_per_step": e2e["prefetch"]["ms_per_step"],
               "detections_per_step": e2e["detections_per_step"], "api": e2e["prefetch"]["api"],
               "note": "every step copies one full batch from pinned host memory (H2D, overlapped with the previous step's "
                       "forward by the copy engine) and reads the detections back (D2H) + host NMS",
               "serial": serial, "u8_input": e2e["u8_input"], "dense_heads": dense}

        # ---- roofline of the dominant kernel (conv_fprop_kernel), rank 0 ---------------------------------
        if rank == 0:
            nl = net.n
            ms_layer = np.zeros(nl + 1, np.float32)
            L.check(L.lib().yc_network_profile(net.h, C.c_void_p(x_dev.data_ptr()), 3, ms_layer.ctypes.data_as(C.c_void_p)),
                    "yc_network_profile")
            conv_ms_br = float(sum(ms_layer[i] for i in range(nl) if net.layer_info(i, 0) == 0))
            other_ms_br = float(ms_layer.sum() - conv_ms_br)
            # the step's conv launches (and, separately, its other kernels) replayed back to back as a CUDA graph, as they
            # run inside the timed step: mean of 10 replays, CUDA events on the launch stream
            t_ms, n_l = C.c_float(0), C.c_int(0)
            L.check(L.lib().yc_network_time_ops(net.h, C.c_void_p(x_dev.data_ptr()), 0, 10, C.byref(t_ms), C.byref(n_l)),
                    "yc_network_time_ops(conv)")
            conv_ms, conv_launches_graph = float(t_ms.value), int(n_l.value)
            L.check(L.lib().yc_network_time_ops(net.h, C.c_void_p(x_dev.data_ptr()), 1, 10, C.byref(t_ms), C.byref(n_l)),
                    "yc_network_time_ops(other)")
            other_ms = float(t_ms.value)
            peaks = _peaks()
            ach = flops_img * B / (conv_ms / 1e3) / 1e12
            # conv LAUNCHES per step: a 1x1 conv computed by its CSP sibling's kernel (fused pair) has none of its own
            n_conv = sum(1 for i in range(nl) if net.layer_info(i, 0) == 0 and net.layer_info(i, 14) < 0)
            roof = {"bound": "tensor", "kernel": "conv_fprop_kernel", "achieved": ach, "peak": peaks["sustained"],
                    "unit": "TFLOP/s", "frac": ach / peaks["sustained"], "traffic": None,
                    "frac_of_burst_peak": ach / peaks["burst"], "peak_source": peaks["source"] + ", sustained bf16",
                    "launches_per_step": n_conv, "flops_per_step": flops_img * B, "conv_ms_per_step": conv_ms,
                    "other_kernels_ms_per_step": other_ms,
                    "avg_launch_us": 1e3 * conv_ms / n_conv,
                    "whole_step_frac": (flops_img * B / (ms_per_step / 1e3) / 1e12) / peaks["sustained"],
                    "conv_ms_per_step_bracketed": conv_ms_br, "other_kernels_ms_per_step_bracketed": other_ms_br,
                    "achieved_bracketed": flops_img * B / (conv_ms_br / 1e3) / 1e12,
                    "how": "sum of algorithmic FLOPs of the conv launches of one step / their duration: the step's "
                           f"{conv_launches_graph} conv launches replayed back to back as one CUDA graph (yc_network_time_ops, mean of 10 "
                           "replays, CUDA events on the launch stream) = how they run inside the timed step; *_bracketed = every "
                           "launch between its own pair of events (yc_network_profile, 3 passes): each kernel then starts on "
                           "an idle GPU, which costs the ~60 short layers several us each"}
            tr = _ncu_traffic(n_conv)
            if tr is not None:
                # compulsory bytes of one step: every conv reads its input and writes its output once (bf16), weights once
                alg = 0.0
                for i in range(nl):
                    if net.layer_info(i, 0) == 0:
                        hw_out = net.layer_info(i, 2) * net.layer_info(i, 3)
                        st_ = max(1, net.layer_info(i, 7))
                        alg += 2.0 * B * hw_out * (net.layer_info(i, 1) + net.layer_info(i, 5) * st_ * st_)
                        alg += 2.0 * net.layer_info(i, 1) * net.layer_info(i, 5) * net.layer_info(i, 6) ** 2
                roof["traffic"] = tr["bytes_per_launch"]
                roof["traffic_unit"] = "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, mean of the step's conv launches)"
                roof["traffic_source"] = tr["source"]
                roof["algorithmic_bytes_per_launch"] = alg / n_conv
            if args.layers:
                for i in range(nl):
                    if ms_layer[i] > 0:
                        t = net.layer_type(i)
                        extra = ""
                        if t == "convolutional":
                            f = 2.0 * net.layer_info(i, 1) * net.layer_info(i, 5) * net.layer_info(i, 6) ** 2 * net.layer_info(i, 2) * net.layer_info(i, 3) * B
                            extra = (f" k{net.layer_info(i,6)} s{net.layer_info(i,7)} {net.layer_info(i,5)}->{net.layer_info(i,1)} "
                                     f"@{net.layer_info(i,2)} bn={net.layer_info(i,10)} st={net.layer_info(i,11)} tiles={net.layer_info(i,12)} "
                                     f"{f/ms_layer[i]/1e9:.0f} TF/s")
                        print(f"  L{i:3d} {t:14s} {ms_layer[i]*1e3:8.1f} us{extra}", file=sys.stderr)
                print(f"  input convert {ms_layer[nl]*1e3:8.1f} us", file=sys.stderr)
            if not args.no_cpu_baseline and world == 1:
                cpu = cpu_baseline(cfgname, hw, args.cpu_budget)

    # ---- data-parallel training (BASELINE configs[3], configs[4]) in the same run: every rank takes part ----------
    train = None
    if not args.quick_exit_after_timed and not args.no_train_leg:
        net.close()
        del net, x_dev
        torch.cuda.empty_cache()
        train = {"yolov4_tiny_416": train_leg(rank, world, dist, "yolov4-tiny", 416, 64, 10),
                 "gaussian_yolov3_608": train_leg(rank, world, dist, "gaussian-yolov3", 608, 8, 10)}
    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
               "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "bf16", "data": "synthetic",
               "config": {"workload": wl, "net": cfgname, "width": hw, "height": hw, "batch_per_gpu": B,
                          "global_batch": B * world, "parallelism": f"image-sharded x{world}, no collective",
                          "weights": "random-init (reference init law), BN folded", "conv_gflop_per_image": flops_img / 1e9,
                          "l2_policy": "per-step working set (>20 GB of activations at batch 64) far exceeds the 126 MB L2; "
                                       "no explicit flush"},
               "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "kernels_per_step": kernels_per_forward,
               "roofline": roof, "cpu_baseline": cpu, "train": train}
        emit(out)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
