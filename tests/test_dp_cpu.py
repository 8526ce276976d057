"""Host-side logic of the data-parallel path on CPU: world_size-2 gloo group (no GPU, no NCCL): batch sharding
(GetPartialData, data.cpp:890-901), the unique-id hand-off, the bucket layout the all-reduce uses, and — with the
reference CPU path as the arithmetic — that summing per-shard gradients over ranks reproduces what one replica computes
on the concatenated batch when the batchnorm statistics are per shard (the oracle is run per shard and
summed)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from darknet_b200 import cfg as cfgmod, dp, lib as L


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. id hand-off: only rank 0's id survives
        uid = dp.exchange_unique_id(dist, lambda: bytes([(7 * i + 1) % 256 for i in range(128)]) if rank == 0 else b"\0" * 128)
        assert uid == bytes([(7 * i + 1) % 256 for i in range(128)])
        # 2. sharding covers the batch exactly once
        x = cfgmod.synth_images(8, 3, 4, 4, seed=3)
        mine = dp.shard(x, rank, world)
        parts = [torch.zeros_like(torch.from_numpy(mine)) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(mine))
        assert np.array_equal(np.concatenate([p.numpy() for p in parts]), x)
        # 3. an all-reduce (sum) of per-rank gradient buckets == the sum computed in one place
        layout = dp.bucket_layout([(0, 16, 3, 3), (1, 32, 16, 3), (2, 18, 32, 1)])
        rng = np.random.default_rng(100 + rank)
        buckets = [torch.from_numpy(rng.standard_normal(n).astype(np.float32)) for _, n in layout]
        want = [sum(np.random.default_rng(100 + r).standard_normal(n).astype(np.float32) for r in range(world)) for _, n in layout]
        # regenerate deterministically per bucket order (same generator sequence as above)
        gens = [np.random.default_rng(100 + r) for r in range(world)]
        want = []
        for _, n in layout:
            want.append(sum(g.standard_normal(n).astype(np.float32) for g in gens))
        for b in buckets:
            dist.all_reduce(b)
        for b, w in zip(buckets, want):
            assert np.allclose(b.numpy(), w, atol=1e-5)
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_dp_host_logic_gloo_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in ps:
        p.join(timeout=30)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def test_bucket_layout_matches_library():
    lib = L.lib()
    for (n, c, k) in [(32, 3, 3), (64, 32, 3), (255, 256, 1), (512, 256, 3), (18, 64, 1)]:
        (_, floats), = dp.bucket_layout([(0, n, c, k)])
        assert floats == lib.yc_conv_wgrad_bytes(n, c, k) // 4 + 2 * ((n + 3) // 4 * 4)


def test_shard_bounds():
    assert dp.shard_bounds(64, 3, 8) == (24, 32)
    with pytest.raises(AssertionError):
        dp.shard_bounds(10, 0, 4)
