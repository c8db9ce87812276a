"""GPU-side init: the wavetables synthesised on the device.

not-gpu: tbf_build_tonegen_tables_synth leaves every table but the waves identical to the full
builder's (the click envelopes' place in the rand() stream included), and tbf_wave_sample /
tbf_dither_bits reproduce the full builder's samples one by one.
gpu: the device's waves are bit-identical to the host-built ones for tunings spread over the
Scala archive and for the reference's osc-*.cfg variants (12 partials, other wave shapes), and
an organ of a synthesised set renders bit-identically to the oracle."""
import ctypes as C

import numpy as np
import pytest

from tunebfree_b200 import cabi, tables, workloads


def _waves(tt):
    return np.ctypeslib.as_array(tt.wave_data, shape=(int(tt.wave_total),)).copy()


def _same_but_waves(a, b):
    for fname, _ in a._fields_:
        va, vb = getattr(a, fname), getattr(b, fname)
        if isinstance(va, C._Pointer):
            continue
        if isinstance(va, C.Array):
            assert bytes(va) == bytes(vb), fname
        else:
            assert va == vb, fname
    nc = a.contrib_start[cabi.MAX_KEYS]
    for name in ("contrib_wheel", "contrib_bus", "contrib_level"):
        pa = np.ctypeslib.as_array(getattr(a, name), shape=(int(nc),))
        pb = np.ctypeslib.as_array(getattr(b, name), shape=(int(nc),))
        assert pa.tobytes() == pb.tobytes(), name


CASES = [("", 1, None), ("", 7, 19), (None, 3, 777)]  # (cfg | None = sawtooth variant, seed, scale index)


def _spec(cfg, seed, scale):
    from tunebfree_b200 import config5
    if cfg is None:
        cfg = config5.OSC_VARIANTS["sawtooth"]
    freqs = None
    if scale is not None:
        freqs = tables.freqs_from_ratios(workloads.scala_sweep()[1][scale])
    return tables.InstanceSpec(freqs=freqs, seed=seed, cfg=cfg)


@pytest.mark.parametrize("cfg,seed,scale", CASES)
def test_synth_builder_equals_full_builder_but_for_the_waves(cfg, seed, scale):
    full = _spec(cfg, seed, scale).build_tonegen()
    part, ws = _spec(cfg, seed, scale).build_tonegen_synth()
    assert not part.wave_data and part.wave_total == full.wave_total
    _same_but_waves(full, part)
    # the host's per-sample routine + the dither stream reproduce the builder's samples
    w = _waves(full)
    n = int(full.wave_total)
    bits = np.zeros((n + 31) // 32, dtype=np.uint32)
    tables._lib().tbf_dither_bits(seed, n, bits.ctypes.data)
    rng = np.random.RandomState(5)
    for wheel in (1, 17, 92, 200, 256):
        off, ln = full.wave_offset[wheel], full.wave_length[wheel]
        for i in [0, 1, ln - 1] + rng.randint(0, ln, size=20).tolist():
            k = off + i
            v = tables._lib().tbf_wave_sample(C.byref(ws), wheel, i, int((bits[k >> 5] >> (k & 31)) & 1))
            assert np.float32(v).view(np.uint32) == w[k].view(np.uint32), (wheel, i)
    tables.free_tables(full, None)
    tables.free_tables(part, None)


@pytest.mark.gpu
def test_device_waves_bit_identical_to_host_waves(ref):
    from tunebfree_b200 import config5, engine
    ctx = engine.Context(0)
    names, degs, _ = workloads.scala_sweep()
    specs = [tables.InstanceSpec(freqs=tables.freqs_from_ratios(degs[g]), seed=1)
             for g in range(0, len(names), len(names) // 40)]
    specs += [tables.InstanceSpec(seed=2 + k, cfg=cfg) for k, cfg in enumerate(config5.OSC_VARIANTS.values())]
    specs += [tables.InstanceSpec(seed=9, rate=r) for r in (22050.0, 44100.0, 96000.0)]
    for sp in specs:
        full = sp.build_tonegen()
        part, ws = sp.build_tonegen_synth()
        h = ctx.tonegen_tables_synth(part, ws)
        got = ctx.read_waves(h, full.wave_total)
        assert np.array_equal(got.view(np.uint32), _waves(full).view(np.uint32))
        ctx.destroy_tonegen_tables(h)
        tables.free_tables(full, None)
        tables.free_tables(part, None)
    flagged, differed = ctx.synth_stats()
    print("wave synthesis: %d table sets, %d samples flagged, %d differed from glibc" % (len(specs), flagged, differed))
    # an organ on a synthesised set against the oracle
    sp = tables.InstanceSpec(freqs=tables.freqs_from_ratios(degs[123]), seed=1)
    part, ws = sp.build_tonegen_synth()
    tt0, wt0 = tables.InstanceSpec(seed=1).build()
    i = ctx.create_instances(ctx.tonegen_tables_synth(part, ws), ctx.whirl_tables(wt0), 1)
    ev = workloads.instance_events(0, 200, config=4)
    gl, gr = ctx.render(1, 200, [ev], cabi.RENDER_WHIRLPROC3, ids=[i])
    r = ref.RefInstance(freqs=tables.freqs_from_ratios(degs[123]), seed=1)
    rl, rr = r.render(200, cabi.RENDER_WHIRLPROC3, ev)
    r.close()
    assert np.array_equal(gl[0].view(np.uint32), rl.view(np.uint32))
    assert np.array_equal(gr[0].view(np.uint32), rr.view(np.uint32))
    tables.free_tables(part, None)
    tables.free_tables(tt0, wt0)
    ctx.close()


@pytest.mark.gpu
def test_organs_create_pipeline_matches_one_by_one(ref):
    """tbf_gpu_organs_create (builder threads + in-order uploads, wavetables on the device) gives
    the organs that building and uploading one by one on the host gives: same waves, and the
    render of a few of them is the oracle's."""
    from tunebfree_b200 import engine
    names, degs, _ = workloads.scala_sweep()
    pick = list(range(5, len(names), len(names) // 60))[:60]
    ctx = engine.Context(0)
    tt0, wt0 = tables.InstanceSpec(seed=1).build()
    wh = ctx.whirl_tables(wt0)
    tables.free_tables(tt0, wt0)
    specs = (cabi.TonegenSpec * len(pick))()
    for k, g in enumerate(pick):
        sp = tables.InstanceSpec(freqs=tables.freqs_from_ratios(degs[g]), seed=1)
        C.memmove(C.byref(specs[k]), C.byref(sp.tg), C.sizeof(cabi.TonegenSpec))
    ids, handles, (wall, wait, up) = ctx.create_organs(specs, wh, threads=4)
    assert len(set(ids.tolist())) == len(pick) and len(set(handles)) == len(pick) and wall > 0
    for k in (0, 17, 59):
        full = tables.InstanceSpec(freqs=tables.freqs_from_ratios(degs[pick[k]]), seed=1).build_tonegen()
        got = ctx.read_waves(handles[k], full.wave_total)
        assert np.array_equal(got.view(np.uint32), _waves(full).view(np.uint32)), names[pick[k]]
        tables.free_tables(full, None)
    n_frag = 120
    ev = workloads.instance_events(0, n_frag, config=4)
    sel = [3, 31, 58]
    gl, gr = ctx.render(len(sel), n_frag, [ev] * len(sel), cabi.RENDER_WHIRLPROC3, ids=[int(ids[k]) for k in sel])
    for j, k in enumerate(sel):
        r = ref.RefInstance(freqs=tables.freqs_from_ratios(degs[pick[k]]), seed=1)
        rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, ev)
        r.close()
        assert np.array_equal(gl[j].view(np.uint32), rl.view(np.uint32)), names[pick[k]]
        assert np.array_equal(gr[j].view(np.uint32), rr.view(np.uint32)), names[pick[k]]
    # a spec the builder refuses stops the pipeline with its error code
    bad = (cabi.TonegenSpec * 3)()
    for k in range(3):
        C.memmove(C.byref(bad[k]), C.byref(specs[k]), C.sizeof(cabi.TonegenSpec))
    bad[1].sample_rate = -1.0
    with pytest.raises(engine.TbfError):
        ctx.create_organs(bad, wh, threads=2)
    ctx.close()
