"""Data-parallel gradient all-reduce check (run under torchrun, one rank per GPU):
   torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py
Every rank trains the same net on ITS shard of the batch.  Checks that after yc_network_forward_backward with the NCCL
all-reduce enabled every rank holds the SUM of the per-rank gradients (gathered with torch.distributed for the check),
that two steps with updates keep the replicas' weights bit-identical, and prints the overlap timing."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from darknet_b200 import cfg as cfgmod, cfgs, lib as L, network as N
from darknet_b200 import dp


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    L.check(L.lib().yc_set_device(local), "yc_set_device")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    name = sys.argv[1] if len(sys.argv) > 1 else "micro"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    hw = 64 if name == "micro" else 416
    classes = 3 if name == "micro" else 80
    kw = {} if name == "micro" else dict(w=hw, h=hw)
    text = cfgs.CONFIGS[name](batch=B, subdivisions=1, **kw)
    weights = cfgmod.synth_weights(text)
    flags = L.NET_DEFAULT | L.NET_TRAIN | L.NET_TRAIN_REF_CPU

    def fresh():
        return N.LoadSynthetic(text, weights, batch=B, flags=flags)

    # the global batch, sharded by rank (GetPartialData, data.cpp:890-901)
    xg = cfgmod.synth_images(B * world, 3, hw, hw, seed=5)
    tg = cfgmod.synth_truth(B * world, 90, classes=classes, seed=6)
    x, t = dp.shard(xg, rank, world), dp.shard(tg, rank, world)

    # 1. local gradients without communication
    net0 = fresh()
    N.TrainNetworkDatum(net0, x, t)
    convs = [i for i in range(net0.n) if net0.layer_type(i) == "convolutional"]
    local_g = {i: torch.from_numpy(N.layer_param(net0, i, "weight_updates")).cuda() for i in convs}
    net0.close()
    # 2. with the all-reduce
    net = fresh()
    dp.init(net, dist)
    cost = N.TrainNetworkDatum(net, x, t)
    # Since round 2 the training step is bit-reproducible (no float atomics), so the all-reduced bucket equals the sum of
    # the independent local runs up to the order of the NCCL sum (measured: rel-L2 0.0 on two ranks); every rank must hold
    # bit-identical reduced gradients.
    worst, per_layer, identical = 0.0, {}, True
    for i in convs:
        want = local_g[i].clone()
        dist.all_reduce(want)
        got = torch.from_numpy(N.layer_param(net, i, "weight_updates")).cuda()
        err = float((got - want).norm() / (want.norm() + 1e-30))
        solo = float((got - local_g[i]).norm() / (want.norm() + 1e-30))   # would be ~0 if nothing had been reduced
        worst = max(worst, err)
        per_layer[i] = (err, solo)
        g0 = got.clone()
        dist.broadcast(g0, 0)
        identical = identical and bool(torch.equal(got, g0))
    # 2b. the exact statement about the SUM: same run, gradients pulled before and after a deferred reduction
    net2 = fresh()
    dp.init(net2, dist)
    dp.set_overlap(net2, False)
    N.TrainNetworkDatum(net2, x, t)
    keys = lambda i: ["weight_updates", "bias_updates"] + (["scale_updates"] if net2.layer_info(i, 9) else [])
    before = {(i, k): torch.from_numpy(N.layer_param(net2, i, k)).cuda() for i in convs for k in keys(i)}
    dp.reduce(net2)
    exact_err = 0.0
    for (i, k), loc in before.items():
        want = loc.clone()
        dist.all_reduce(want)
        got = torch.from_numpy(N.layer_param(net2, i, k)).cuda()
        exact_err = max(exact_err, float((got - want).abs().max() / (want.abs().max() + 1e-30)))
    # 3. replicas stay identical through updates
    N.UpdateNetwork(net)
    N.TrainNetworkDatum(net, x, t)
    N.UpdateNetwork(net)
    same = True
    for i in convs:
        w = torch.from_numpy(N.layer_param(net, i, "weights")).cuda()
        w0 = w.clone()
        dist.broadcast(w0, 0)
        same = same and bool(torch.equal(w, w0))
    # 4. timing: step with and without the all-reduce (device events, max over ranks)
    xd = torch.from_numpy(x).cuda()
    td = torch.from_numpy(t).cuda()
    st = torch.cuda.Stream()   # explicit stream for the steps and their events (NULL = the network's own stream)
    torch.cuda.synchronize()
    import ctypes as _C
    sp = _C.c_void_p(st.cuda_stream)

    def timed(n, iters=10):
        for _ in range(3):
            if n is net or n is net2:
                dp.discard_step(n)   # timing loop: backward passes without updates
            N.ForwardBackwardNetworkGpu(n, xd.data_ptr(), td.data_ptr(), sp)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(iters):
            if n is net or n is net2:
                dp.discard_step(n)   # timing loop: backward passes without updates
            N.ForwardBackwardNetworkGpu(n, xd.data_ptr(), td.data_ptr(), sp)
        e1.record(st)
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    ms_dp = timed(net)

    def timed_deferred(iters=10):
        def one():
            dp.discard_step(net2)
            N.ForwardBackwardNetworkGpu(net2, xd.data_ptr(), td.data_ptr(), sp)
            dp.reduce(net2, sp)
        for _ in range(3):
            one()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(iters):
            one()
        e1.record(st)
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    ms_deferred = timed_deferred()
    net1 = fresh()
    ms_local = timed(net1)
    # 5. the library refuses to all-reduce a bucket twice between two updates (the momentum term is carried divided by the
    #    world size): a second backward pass under per-layer overlap must fail on every rank alike, before any NCCL call
    import ctypes as _C2
    torch.cuda.synchronize(); dist.barrier()
    dp.discard_step(net)
    args = (net.h, _C2.c_void_p(xd.data_ptr()), _C2.c_void_p(td.data_ptr()), sp)
    rc1 = L.lib().yc_network_forward_backward(*args)
    rc2 = L.lib().yc_network_forward_backward(*args)
    msg = L.lib().yc_last_error().decode()
    guard_ok = rc1 == 0 and rc2 != 0 and "reduced twice" in msg
    N.UpdateNetwork(net, stream=sp)            # consumes the first pass: the next step is legal again
    rc3 = L.lib().yc_network_forward_backward(*args)
    N.UpdateNetwork(net, stream=sp)
    torch.cuda.synchronize(); dist.barrier()
    guard_ok = guard_ok and rc3 == 0
    if rank == 0:
        print(json.dumps({"dp_check": name, "world": world, "batch_per_gpu": B, "cost_rank0": cost,
                          "allreduce_relL2_vs_sum_of_local_runs": worst, "replicas_identical_after_2_updates": same,
                          "deferred_reduce_max_rel_err_vs_torch_allreduce": exact_err,
                          "ms_step_with_allreduce": ms_dp, "ms_step_deferred_allreduce": ms_deferred,
                          "ms_step_local_only": ms_local,
                          "buckets": L.lib().yc_network_dp_info(net.h, 2), "mib_per_step": L.lib().yc_network_dp_info(net.h, 3),
                          "nccl": L.lib().yc_dp_nccl_version(),
                          "reduced_gradients_bit_identical_across_ranks": identical,
                          "second_backward_before_update_is_refused": guard_ok,
                          "per_layer_relL2_vs_sum_and_vs_local_only": {str(k): [float(f"{v[0]:.2e}"), float(f"{v[1]:.2e}")]
                                                                       for k, v in per_layer.items()}}))
    dist.barrier()
    net.close(); net1.close(); net2.close()
    dist.destroy_process_group()
    return 0 if (exact_err < 1e-6 and same and identical and guard_ok) else 1


if __name__ == "__main__":
    sys.exit(main())
