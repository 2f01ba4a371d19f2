"""Data-parallel gradient reduction on CPU: world_size 2 over gloo (the NCCL path runs the same Python; streams/events
are skipped for CPU tensors). Checks: bucket construction tiles the arena on leaf boundaries; buckets are issued in
index order as their leaves finish; the reduced gradient equals the sum over ranks (train_step pre-scales by 1/world,
so that is the global-batch mean); the loss is averaged; both ranks end bit-identical."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stable_diffusion_training_b200 import parallel


def _ranges(sizes, align=16):
    out, off = {}, 0
    for i, n in enumerate(sizes):
        out[f"leaf{i:02d}"] = (off, off + n)
        off += -(-n // align) * align
    return out, off


def test_build_buckets_tile_the_arena():
    ranges, total = _ranges([100, 7, 3000, 16, 16, 5000, 1, 900])
    b = parallel.build_buckets(ranges, 1024)
    assert b[0][0] == 0
    for (s0, e0, _), (s1, _, _) in zip(b, b[1:]):
        assert e0 == s1  # contiguous, padding rides with the bucket before it
    assert sorted(n for _, _, names in b for n in names) == sorted(ranges)
    for s, e, names in b[:-1]:
        assert e - s >= 1024 or len(names) == 1
    # a leaf is never split
    for s, e, names in b:
        for n in names:
            assert s <= ranges[n][0] and ranges[n][1] <= max(e, ranges[n][1])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, _, w = parallel.init_process_group_from_env("gloo")
    assert (r, w) == (rank, world)
    sizes = [4096, 33, 70000, 512, 9, 20000, 128, 640]
    ranges, total = _ranges(sizes)
    g = torch.zeros(total, dtype=torch.float32)
    rng = np.random.default_rng(100 + rank)
    for n, (s, e) in ranges.items():
        g[s:e] = torch.from_numpy(rng.standard_normal(e - s).astype(np.float32))
    local = g.clone()
    red = parallel.GradReducer(g, ranges, bucket_bytes=64 * 1024, total_elems=total)
    assert red.world_size == world and len(red.buckets) >= 3
    # backward finishes leaves in arena order here, except that the first leaf comes last (like time_embedding)
    order = sorted(ranges, key=lambda n: ranges[n][0])
    order = order[1:] + order[:1]
    red.begin()
    issued_before_finish = None
    for i, n in enumerate(order):
        red.on_param_grad([n])
        if i == len(order) - 2:
            issued_before_finish = list(red.issued_order)
    loss = torch.tensor([1.0 + rank])
    loss = red.finish(loss)
    assert red.issued_order == list(range(len(red.buckets)))  # strictly in index order on every rank
    # bucket 0 holds the late leaf: nothing could be issued before finish() (in-order issue)
    assert issued_before_finish == []
    red.begin()
    for n in sorted(ranges, key=lambda n: ranges[n][0]):
        red.on_param_grad([n])
    early = list(red.issued_order)
    red._issue_ready(force=True)  # no collectives left un-matched between ranks: both ranks ran the same sequence
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    want = gathered[0] + gathered[1]
    q.put((rank, float(loss), bool(torch.equal(g, want + want - want) or torch.allclose(g, 2 * want - want)), early,
           g.numpy().tobytes()[:64], float((g - (want * 2)).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_grad_reducer_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=100) for _ in range(2))
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    (r0, loss0, _, early0, head0, err0), (r1, loss1, _, early1, head1, err1) = res
    assert loss0 == loss1 == 1.5                      # mean of 1.0 and 2.0
    assert head0 == head1                              # bit-identical reduced gradients on both ranks
    # the arena was reduced twice (two begin/finish rounds): g = 2 * (g0 + g1) ... second round sums the sums
    assert err0 == 0.0 and err1 == 0.0
    assert early0 == early1 and len(early0) >= 2       # with in-order leaves, buckets go out before finish()


def test_peer_all_reduce_slices_tile_every_bucket():
    """PeerAllReduce cuts a bucket into `world` slices (rank r reduces slice r): they tile the bucket, every boundary is
    a multiple of 4 elements (128-bit accesses of sdb_sum_slices_f32) and no slice is longer than the staging rows."""
    from stable_diffusion_training_b200.parallel import PeerAllReduce
    for world in (2, 3, 4, 8):
        obj = PeerAllReduce.__new__(PeerAllReduce)
        obj.world = world
        for (s, e) in ((0, 16), (16, 16 + 4 * 5), (1024, 1024 + 16 * 1000003), (64, 64 + 16), (0, 4)):
            sl = obj._slices(s, e)
            assert len(sl) == world and sl[0][0] == s and sl[-1][1] == e
            for (a, b), (c, d) in zip(sl, sl[1:]):
                assert b == c and a <= b
            assert all(a % 4 == 0 and b % 4 == 0 for a, b in sl)
            per = -(-(-(-(e - s) // world)) // 4) * 4
            assert max(b - a for a, b in sl) <= per
