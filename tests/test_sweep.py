"""Host logic of the dataset sweep: buckets, rank sharding, the statistics all-reduce (gloo, world_size 2)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mel_oracle as orc
from xtts_b200.sweep import (Bucket, allreduce_stats, frame_count, make_buckets, norms_from_stats, shard_buckets)


def _lengths(n, seed=1234):
    rng = np.random.default_rng(seed)
    return [int(round(22050 * u)) for u in rng.uniform(1.0, 30.0, size=n)]


def test_frame_count_matches_oracle():
    for L in (513, 1024, 22050, 220500, 661500, 2646000):
        assert frame_count(L) == orc.frame_count(L)
    assert frame_count(24000, pad=384) == orc.frame_count(24000, orc.V2_SAME)


def test_buckets_partition_and_bound():
    lens = _lengths(256)
    buckets = make_buckets(lens, 2 * 22050, max_clips=64)
    seen = sorted(i for b in buckets for i in b.indices)
    assert seen == list(range(256))                                  # every clip exactly once
    for b in buckets:
        ls = [lens[i] for i in b.indices]
        assert max(ls) == b.max_len and len(ls) <= 64
        assert max(ls) // (2 * 22050) == min(ls) // (2 * 22050)      # one 2 s band per bucket
        assert b.frames == sum(orc.frame_count(L) for L in ls)
        assert b.max_len * len(ls) <= 64 * 220500
    assert len({b.max_len // (2 * 22050) for b in buckets}) <= 15    # <= 15 bands for 1..30 s
    waste = sum(b.max_len * len(b.indices) for b in buckets) / sum(lens)
    assert waste < 1.08                                              # padding overhead of the buckets


def test_empty_and_single():
    assert make_buckets([], 1000) == []
    b = make_buckets([700], 1000)
    assert b == [Bucket((0,), 700, orc.frame_count(700))]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_sharding_balances_frames(world):
    lens = _lengths(2000, seed=4321)
    buckets = make_buckets(lens, 2 * 22050, max_clips=64)
    owned = shard_buckets(buckets, world)
    flat = sorted(j for o in owned for j in o)
    assert flat == list(range(len(buckets)))                         # disjoint cover
    loads = [sum(buckets[j].frames for j in o) for o in owned]
    assert max(loads) - min(loads) <= max(b.frames for b in buckets) # within one batch
    assert max(loads) / (sum(loads) / world) < 1.05


def test_norms_from_stats():
    lm = np.random.default_rng(0).standard_normal((3, 80, 50))
    acc = torch.from_numpy(orc.mel_stats(lm, [50, 20, 7]))
    mean, std = norms_from_stats(acc)
    assert np.allclose(mean.numpy(), orc.norms_from_stats(acc.numpy()), rtol=1e-6)
    allv = np.concatenate([lm[0], lm[1][:, :20], lm[2][:, :7]], axis=1)
    assert np.allclose(std.numpy(), allv.std(axis=1), rtol=1e-5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, lens, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        buckets = make_buckets(lens, 2 * 22050, max_clips=8)
        owned = shard_buckets(buckets, world)
        acc = torch.zeros(161, dtype=torch.float64)
        for j in owned[rank]:                      # the oracle stands in for the GPU kernels on this CPU box
            for i in buckets[j].indices:
                x = orc.synth("uniform", 1, lens[i], 9000 + i)
                lm = orc.callsite_logmel(x, orc.V1, None).numpy()
                acc += torch.from_numpy(orc.mel_stats(lm, [lm.shape[2]]))
        allreduce_stats(acc)
        ret[rank] = acc.numpy()
    finally:
        dist.destroy_process_group()


def test_two_rank_allreduce_equals_single_rank():
    lens = [int(v) for v in np.random.default_rng(5).integers(600, 9000, size=24)]
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), lens, ret), nprocs=world, join=True)
    single = np.zeros(161)
    for i, L in enumerate(lens):
        lm = orc.callsite_logmel(orc.synth("uniform", 1, L, 9000 + i), orc.V1, None).numpy()
        single += orc.mel_stats(lm, [lm.shape[2]])
    assert np.array_equal(ret[0], ret[1])                               # every rank derives identical norms
    assert ret[0][160] == single[160] == sum(orc.frame_count(L) for L in lens)
    assert np.allclose(ret[0], single, rtol=1e-12, atol=0)              # fp64 all-reduce == single-rank result
