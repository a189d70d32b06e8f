"""Host-side multi-GPU logic on CPU: batch sharding and the loss-scalar all-reduce (gloo, world_size 2)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from smplb200 import dist as sd


def test_shard_ranges_partition_the_batch():
    for n in (0, 1, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            rs = [sd.shard_range(n, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
            sizes = sd.shard_sizes(n, world)
            assert max(sizes) - min(sizes) <= 1 and sum(sizes) == n


def _worker(rank, world, port, n, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    # every rank owns a shard of per-sample squared errors; its loss share is divided by the GLOBAL count
    rng = np.random.default_rng(0)
    err = rng.normal(size=(n, 24, 2))
    a, b = sd.shard_range(n, rank, world)
    share = torch.tensor([0.024 * float((err[a:b] ** 2).sum()) / (n * 48), 0.0, 1.0], dtype=torch.float64)
    g = sd.global_loss(share)
    # the side-stream helper (CPU path: synchronous) gives the same sum, slot after slot
    red = sd.SideStreamAllReduce(1, 'cpu')
    for k in range(3):
        red.push(share[:1].float() * (k + 1))
    assert abs(float(red.latest()[0]) - 3 * float(g[0])) <= 1e-5 * abs(float(g[0]))
    if rank == 0:
        np.save(out, np.array([float(g[0]), 0.024 * float((err ** 2).mean())]))
    dist.barrier()
    dist.destroy_process_group()


def test_loss_allreduce_gloo_world2(tmp_path):
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / 'res.npy')
    mp.spawn(_worker, args=(2, port, 101, out), nprocs=2, join=True)
    got, ref = np.load(out)
    assert abs(got - ref) <= 1e-12 * abs(ref)


def test_allreduce_is_noop_without_group():
    t = torch.tensor([1.5, 2.5])
    assert torch.equal(sd.allreduce_scalars(t.clone()), t)
