"""N > 1 host logic on CPU: two gloo ranks shard a batch of organs, no data-path collective."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, per_gpu, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from tunebfree_b200 import sharding, workloads
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    first, per_set = sharding.shard_plan(rank, world, per_gpu, 16)
    handles = []

    def create(s, count):  # stands in for tbf_gpu_instance_create: handles grow per context
        h = sum(c for _, c in handles)
        handles.append((s, count))
        return h

    ids = sharding.slot_instance_ids(per_set, create)
    ev = workloads.batch_events_numpy(first, per_gpu, 40, config=3)
    # every rank's global organ ids, gathered only for the check (not on the data path)
    mine = torch.arange(first, first + per_gpu, dtype=torch.int64)
    got = [torch.zeros(per_gpu, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(got, mine)
    rate, ms = sharding.whole_job_rate(per_gpu * 10.0, 100.0 * (rank + 1), dist)
    q.put((rank, first, ids.tolist(), ev[:, 1:].tolist()[:50], torch.cat(got).tolist(), rate, ms))
    dist.destroy_process_group()


def test_two_ranks_disjoint_cover_and_max_time():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, world, per_gpu = _free_port(), 2, 48
    procs = [ctx.Process(target=_worker, args=(r, world, port, per_gpu, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, f0, ids0, ev0, all0, rate0, ms0), (r1, f1, ids1, ev1, all1, rate1, ms1) = res
    assert (f0, f1) == (0, per_gpu)
    assert all0 == all1 == list(range(world * per_gpu))          # disjoint cover of the job
    assert sorted(ids0) == list(range(per_gpu)) and sorted(ids1) == list(range(per_gpu))
    assert ev0 != ev1                                             # different organs play different chords
    # slowest rank (200 ms) sets the job's time; value = units of ALL ranks / that time
    assert ms0 == ms1 == 200.0 and rate0 == rate1 == pytest.approx(2 * per_gpu * 10.0 / 0.2)


def test_shard_plan_matches_single_process_generation():
    sys.path.insert(0, ROOT)
    from tunebfree_b200 import sharding, workloads
    whole = workloads.batch_events_numpy(0, 8, 30, config=3)
    for rank in range(2):
        first, per_set = sharding.shard_plan(rank, 2, 4, 16)
        part = workloads.batch_events_numpy(first, 4, 30, config=3)
        ref = whole[(whole[:, 0] >= first) & (whole[:, 0] < first + 4)].copy()
        ref[:, 0] -= first
        assert np.array_equal(part, ref)
        assert sorted(s for slots in per_set for s in slots) == [0, 1, 2, 3]
        assert all((first + s) % 16 == k for k, slots in enumerate(per_set) for s in slots)
    with pytest.raises(ValueError):
        sharding.shard_plan(2, 2, 4, 16)


def test_grouped_plan_sorts_by_table_key_and_partitions_every_organ():
    """shard_plan_grouped: organs sharing a table key are neighbours, each GPU gets one block of
    the sorted list, every organ is played exactly once."""
    from tunebfree_b200 import sharding, workloads
    key = lambda g: workloads.organ_variant(g, 5)
    seen = []
    for rank in range(2):
        gids, groups = sharding.shard_plan_grouped(rank, 2, 480, key)
        assert len(gids) == 480 and sum(len(v) for v in groups.values()) == 480
        keys = [key(int(g)) for g in gids]
        assert keys == sorted(keys)
        for k, slots in groups.items():
            assert slots == list(range(slots[0], slots[0] + len(slots)))
            assert all(key(int(gids[s])) == k for s in slots)
        ids = sharding.slot_instance_ids(groups, lambda k, count: 1000 * len(seen) + 0)
        assert len(ids) == 480
        seen += gids.tolist()
    assert sorted(seen) == list(range(960))
    # all 240 (osc, whirl, seed) triples occur, 30 cfg pairs x 16 seeds reached through lcm(30, 16)
    assert len({key(g) for g in range(960)}) == 240
    assert workloads.organ_variant(7, 3) == ("default", "default", 8)
    with pytest.raises(ValueError):
        sharding.shard_plan_grouped(3, 2, 4, key)


def test_sweep_plan_covers_every_scale_once():
    """Config 4: the ranks' blocks of the Scala sweep are disjoint and cover it; a rank asked for
    the whole sweep gets every scale once, rotated by its rank."""
    from tunebfree_b200 import sharding, workloads
    n = len(workloads.scala_sweep()[0])
    assert n == 5176
    per = n // 8
    seen = np.concatenate([sharding.shard_plan_sweep(r, 8, per, n) for r in range(8)])
    assert len(np.unique(seen)) == 8 * per == n
    whole = sharding.shard_plan_sweep(3, 8, n, n)
    assert sorted(whole.tolist()) == list(range(n)) and whole[0] == (3 * n) % n
    with pytest.raises(ValueError):
        sharding.shard_plan_sweep(8, 8, 10, n)
