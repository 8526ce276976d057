"""Host-side plumbing for data-parallel training (one process per GPU): batch sharding and the NCCL unique-id
exchange over torch.distributed.  The collective itself runs inside libyc_b200.so (csrc/network_dp.cu).

Replaces the host side of the reference's TrainNetworks (network_kernels.cu:405-484: one pthread per GPU) and
GetPartialData (data.cpp:890-901)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as L


def shard_bounds(n: int, rank: int, world: int):
    """Rows [lo, hi) of a global batch of n that belong to `rank` (GetPartialData: contiguous equal parts)."""
    assert n % world == 0, "global batch must divide evenly over the replicas (data.cpp:892)"
    per = n // world
    return rank * per, (rank + 1) * per


def shard(a: np.ndarray, rank: int, world: int) -> np.ndarray:
    lo, hi = shard_bounds(a.shape[0], rank, world)
    return np.ascontiguousarray(a[lo:hi])


def bucket_layout(layers):
    """[(layer, floats)] of the per-layer gradient buckets [weight_updates | bias_updates | scale_updates] in the order
    the backward pass all-reduces them (last conv first).  `layers` = [(index, out_c, in_c, ksize)]."""
    out = []
    for idx, n, c, k in reversed(list(layers)):
        ldw = (c + 31) // 32 * 32
        out.append((idx, n * k * k * ldw + 2 * ((n + 3) // 4 * 4)))
    return out


def exchange_unique_id(dist, make_id):
    """Rank 0 creates the 128-byte NCCL id with `make_id()`; everyone receives it through torch.distributed."""
    import torch
    rank = dist.get_rank()
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    buf = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        buf.copy_(torch.frombuffer(bytearray(make_id()), dtype=torch.uint8))
    dist.broadcast(buf, 0)
    return bytes(buf.cpu().numpy().tobytes())


def _make_id():
    raw = (C.c_ubyte * 128)()
    L.check(L.lib().yc_dp_unique_id(raw), "yc_dp_unique_id")
    return bytes(raw)


def init(net, dist):
    """Join `net` (loaded with NET_TRAIN) to the data-parallel world of the initialised torch.distributed group."""
    uid = exchange_unique_id(dist, _make_id)
    raw = (C.c_ubyte * 128).from_buffer_copy(uid)
    L.check(L.lib().yc_network_dp_init(net.h, raw, dist.get_rank(), dist.get_world_size()), "yc_network_dp_init")


def set_overlap(net, on: bool):
    """Per-layer all-reduce overlapped with backward (default, subdivisions=1) or one grouped reduction before the update."""
    L.check(L.lib().yc_network_dp_set_overlap(net.h, 1 if on else 0), "yc_network_dp_set_overlap")


def reduce(net, stream=None):
    """All-reduce the gradients accumulated since the last reduction (collective; no-op if none are pending)."""
    L.check(L.lib().yc_network_dp_reduce(net.h, stream), "yc_network_dp_reduce")


def discard_step(net):
    """Timing loops that run backward passes without updates: declare this step's gradients discarded, so that the next
    backward pass may all-reduce again (the library refuses to reduce a bucket twice between two updates)."""
    L.check(L.lib().yc_network_dp_discard_step(net.h), "yc_network_dp_discard_step")
