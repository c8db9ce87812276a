"""Multi-GPU sharding of a batch of organs: one process per GPU, no data-path collective.

Organs are independent: rank r renders global organs
[r * per_gpu, (r + 1) * per_gpu), each with the table set `global id mod n_table_sets`.
torch.distributed is used only around the timed region (barrier, max of the device times)."""
import numpy as np


def shard_plan(rank, world, per_gpu, n_table_sets):
    """-> (first_global_id, per_set) where per_set[s] lists the batch slots of this rank that
    play with table set s (slot i plays global organ first_global_id + i)."""
    if not (0 <= rank < world) or per_gpu <= 0 or n_table_sets <= 0:
        raise ValueError("bad shard arguments")
    first = rank * per_gpu
    per_set = [[] for _ in range(n_table_sets)]
    for slot in range(per_gpu):
        per_set[(first + slot) % n_table_sets].append(slot)
    return first, per_set


def shard_plan_grouped(rank, world, per_gpu, key_of):
    """Block partition of the organ list SORTED BY TABLE KEY: organs that share
    wave tables become neighbours in the batch and land on one GPU, so the warps in flight
    at any time read a few table sets out of L2 instead of all of them out of HBM.
    key_of(global_id) -> hashable, sortable.  -> (global ids of this rank's slots,
    {key: [slots]} in slot order)."""
    if not (0 <= rank < world) or per_gpu <= 0:
        raise ValueError("bad shard arguments")
    order = sorted(range(world * per_gpu), key=lambda g: (key_of(g), g))
    mine = order[rank * per_gpu:(rank + 1) * per_gpu]
    groups = {}
    for slot, g in enumerate(mine):
        groups.setdefault(key_of(g), []).append(slot)
    return np.asarray(mine, dtype=np.int64), groups


def shard_plan_sweep(rank, world, per_gpu, n_items):
    """Config 4 (one organ per Scala scale): rank r plays items [r * per_gpu, (r + 1) * per_gpu)
    of the sweep, wrapping around its end, so with per_gpu * world == n_items every scale is
    rendered exactly once and with per_gpu == n_items every GPU renders the whole sweep (weak
    scaling).  -> int64 item index of every batch slot."""
    if not (0 <= rank < world) or per_gpu <= 0 or n_items <= 0:
        raise ValueError("bad shard arguments")
    return (rank * per_gpu + np.arange(per_gpu, dtype=np.int64)) % n_items


def slot_instance_ids(per_set, create):
    """create(set index or key, count) -> first instance handle of `count` new instances.
    per_set: list indexed by set (shard_plan) or {key: slots} (shard_plan_grouped).
    Returns the uint32 handle of every batch slot."""
    items = list(per_set.items()) if isinstance(per_set, dict) else list(enumerate(per_set))
    n = sum(len(slots) for _, slots in items)
    ids = np.zeros(n, dtype=np.uint32)
    for s, slots in items:
        if slots:
            first = create(s, len(slots))
            ids[slots] = np.arange(first, first + len(slots), dtype=np.uint32)
    return ids


def whole_job_rate(organ_seconds_per_rank, local_ms, dist=None, device="cpu"):
    """Whole-job organ-seconds per second: units all ranks rendered / slowest rank's time."""
    world = 1
    ms = float(local_ms)
    if dist is not None:
        import torch
        world = dist.get_world_size()
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return world * organ_seconds_per_rank / (ms * 1e-3), ms
