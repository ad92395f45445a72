"""CPU, world_size 2 over gloo: the host-side logic of the N > 1 paths (robot partition, sample-shard slices, hand-out
of the ncclUniqueId payload).  The device-side merge of shard records is covered by tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from quadruped_pympc_b200.distributed import partition_robots, shard_slice, broadcast_bytes, all_gather_bytes


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, R, N, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        lo, hi = partition_robots(R, rank, world)
        b, e = shard_slice(N, rank, world)
        mine = torch.tensor([lo, hi, b, e], dtype=torch.int64)
        allr = [torch.zeros(4, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allr, mine)
        payload = bytes(range(128)) if rank == 0 else None
        got = broadcast_bytes(payload, 128, 0)
        # hand-out of the peer-exchange mailbox descriptors: every rank contributes 128 bytes, all get them in rank order
        blobs = all_gather_bytes(bytes([rank]) * 128)
        assert blobs == [bytes([r]) * 128 for r in range(world)]
        # robots partition with no communication: each rank reports only a count (ticks/s is summed by the launcher)
        ticks = torch.tensor([float(hi - lo)])
        dist.all_reduce(ticks)
        if rank == 0:
            out.put(([t.tolist() for t in allr], got == bytes(range(128)), float(ticks.item())))
        else:
            assert got == bytes(range(128))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('R,N', [(4096, 10000), (7, 1001), (1, 3)])
def test_partitions_cover_everything_exactly_once(R, N):
    world = 2
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, R, N, q)) for r in range(world)]
    for p in procs:
        p.start()
    slices, id_ok, total = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert id_ok and total == R
    robots = sorted((s[0], s[1]) for s in slices)
    samples = sorted((s[2], s[3]) for s in slices)
    assert robots[0][0] == 0 and robots[-1][1] == R and all(a[1] == b[0] for a, b in zip(robots, robots[1:]))
    assert samples[0][0] == 0 and samples[-1][1] == N and all(a[1] == b[0] for a, b in zip(samples, samples[1:]))


def test_slices_for_eight_ranks():
    for n in (1000, 10000, 1_000_000, 1337):
        cover = [shard_slice(n, r, 8) for r in range(8)]
        assert cover[0][0] == 0 and cover[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(cover, cover[1:]))
    assert [partition_robots(4096, r, 8) for r in range(8)] == [(512 * r, 512 * (r + 1)) for r in range(8)]
