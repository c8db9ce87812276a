"""Full-size runs (BASELINE.json configs 2, 3, 4 and 5 instance counts) checked through properties
that do not need the audio on the host: a position-keyed checksum of every organ's L/R bit
patterns is formed on the device; sampled organs are compared with the oracle's render of the
same events through the same checksum ("a checksum of checksums"), the run is repeated after a
reset (determinism over the whole batch), and no organ may raise an error flag."""
import numpy as np
import pytest

from tunebfree_b200 import cabi, engine, workloads

pytestmark = pytest.mark.gpu


def _run(ref, ctx, n, n_frag, config, sampled):
    r0 = ref.RefInstance(seed=1)
    tg = ctx.tonegen_tables(r0.tonegen_tables())
    wh = ctx.whirl_tables(r0.whirl_tables())
    r0.close()
    first = ctx.create_instances(tg, wh, n)
    ids = list(range(first, first + n))
    ev_np = workloads.batch_events_numpy(0, n, n_frag, config)
    ev = engine.events_from_numpy(ev_np)
    flags = workloads.render_flags(config)
    cs = np.zeros(n, dtype=np.uint64)
    ctx.render_device(n, n_frag, ev, flags, None, None, n_frag * 128, ids=ids, checksums=cs)
    assert len(set(cs.tolist())) > n * 0.99          # different chords -> different audio
    for i in sampled:
        r = ref.RefInstance(seed=1)
        l, rr = r.render(n_frag, flags, workloads.instance_events(i, n_frag, config))
        r.close()
        assert np.abs(l).max() > 0.01
        assert engine.checksum_host(l, rr) == int(cs[i]), i
    for i in (0, n // 2, n - 1):
        assert ctx.export_state(first + i).error_flags == 0
    ctx.reset(first, n)
    cs2 = np.zeros(n, dtype=np.uint64)
    ctx.render_device(n, n_frag, ev, flags, None, None, n_frag * 128, ids=ids, checksums=cs2)
    assert np.array_equal(cs, cs2)


def test_config3_4096_organs(ref, gpu_ctx):
    """BASELINE config 3 at its full size: 4096 organs x 10 s through tonegen + vibrato + Leslie
    (whirlProc3); 8 sampled organs against the oracle."""
    _run(ref, gpu_ctx, 4096, 3750, 3, sampled=(0, 1, 31, 32, 1000, 2047, 2048, 4095))


def test_config2_1024_organs(ref, gpu_ctx):
    """BASELINE config 2 at its full size: 1024 organs x 10 s, tone generator only."""
    _run(ref, gpu_ctx, 1024, 3750, 2, sampled=(0, 5, 511, 1023))


def test_config5_cfg_mix_16384_organs(ref, gpu_ctx):
    """BASELINE config 5 at its full size (16 384 organs x 10 s): every pair of the reference's 10 osc x
    3 whirl cfg variants with 16 engine seeds (80 wave-table sets, 3 Leslie table sets, tables from
    the product's own builder), organs grouped by table set.  One organ of each of the 30 pairs
    and a few others are rendered by the oracle -- its own cfg parser fed the same text -- and
    compared through the device checksum."""
    from tunebfree_b200 import sharding, tables
    n, n_frag = 16384, 3750
    variant = lambda g: workloads.organ_variant(g, 5)
    gids, groups = sharding.shard_plan_grouped(0, 1, n, variant)
    assert len(groups) == 240 and sorted(gids.tolist()) == list(range(n))
    tgh, whh = {}, {}
    for osc, whirl, seed in groups:
        if (osc, seed) not in tgh or whirl not in whh:
            spec = tables.InstanceSpec(seed=seed, cfg=workloads.variant_cfg(osc, whirl))
            tt, wt = spec.build()
            assert spec.rejected == []
            if (osc, seed) not in tgh:
                tgh[(osc, seed)] = gpu_ctx.tonegen_tables(tt)
            if whirl not in whh:
                whh[whirl] = gpu_ctx.whirl_tables(wt)
            tables.free_tables(tt, wt)
    assert len(tgh) == 80 and len(whh) == 3
    ids = sharding.slot_instance_ids(
        groups, lambda k, count: gpu_ctx.create_instances(tgh[(k[0], k[2])], whh[k[1]], count))
    ev = engine.events_from_numpy(workloads.batch_events_numpy(0, n, n_frag, 5, global_ids=gids))
    flags = workloads.render_flags(5)
    cs = np.zeros(n, dtype=np.uint64)
    gpu_ctx.render_device(n, n_frag, ev, flags, None, None, n_frag * 128, ids=ids, checksums=cs)
    slot_of = {int(g): s for s, g in enumerate(gids)}
    for g in list(range(30)) + [77, 4099, 9001, 16383]:
        osc, whirl, seed = variant(g)
        r = ref.RefInstance(seed=seed, cfg=workloads.variant_cfg(osc, whirl))
        l, rr = r.render(n_frag, flags, workloads.instance_events(g, n_frag, 5))
        r.close()
        assert np.abs(l).max() > 0.01
        assert engine.checksum_host(l, rr) == int(cs[slot_of[g]]), (g, osc, whirl, seed)
    for s in (0, n // 3, n - 1):
        assert gpu_ctx.export_state(int(ids[s])).error_flags == 0


def test_config4_scala_sweep_all_scales(ref):
    """BASELINE config 4 at its full size: one organ per scale of the reference's Scala archive
    (5176 tunings, each with its own wave store built by the product's table builder on the host
    cores), 2 s of the sweep chord through tonegen + vibrato + Leslie.  Every 173rd organ (30) is
    rendered by the oracle from the same 128 frequencies and compared through the device
    checksum; different tunings must give different audio."""
    from concurrent.futures import ThreadPoolExecutor
    import os
    from tunebfree_b200 import tables
    names, degs, _ = workloads.scala_sweep()
    n, n_frag = len(names), 750
    ctx = engine.Context(0)  # own context: 31 GB of wave stores are released with it
    tt0, wt0 = tables.InstanceSpec(seed=1).build()
    wh = ctx.whirl_tables(wt0)
    tables.free_tables(tt0, wt0)

    import ctypes as C
    specs = (cabi.TonegenSpec * n)()
    for g in range(n):
        sp = tables.InstanceSpec(freqs=tables.freqs_from_ratios(degs[g]), seed=1)
        C.memmove(C.byref(specs[g]), C.byref(sp.tg), C.sizeof(cabi.TonegenSpec))
    # builder threads on the host cores, wavetables synthesised on the device (tbf_gpu_organs_create)
    ids, _, _ = ctx.create_organs(specs, wh, threads=os.cpu_count() or 1)
    ids = [int(i) for i in ids]
    flagged, differed = ctx.synth_stats()
    print("config 4: %d organs, %d wave samples flagged, %d repaired with the host's libm" % (n, flagged, differed))
    sc = workloads.instance_events(0, n_frag, config=4)
    ev_np = np.concatenate([np.concatenate([np.full((len(sc), 1), s, dtype=np.int32),
                                            np.array(sorted(sc, key=lambda e: e[0]), dtype=np.int32)], axis=1)
                            for s in range(n)], axis=0)
    ev = engine.events_from_numpy(ev_np)
    flags = workloads.render_flags(4)
    cs = np.zeros(n, dtype=np.uint64)
    ctx.render_device(n, n_frag, ev, flags, None, None, n_frag * 128, ids=ids, checksums=cs)
    # duplicates exist in the archive (the same scale under two names); most tunings differ
    assert len(set(cs.tolist())) > 0.9 * n
    for i in range(0, n, 173):
        r = ref.RefInstance(freqs=tables.freqs_from_ratios(degs[i]), seed=1)
        l, rr = r.render(n_frag, flags, sc)
        r.close()
        assert engine.checksum_host(l, rr) == int(cs[i]), names[i]
    for i in (0, n // 2, n - 1):
        assert ctx.export_state(ids[i]).error_flags == 0
    ctx.close()
