"""Resource lifetime at the boundary and the output options added in round 2.

* tbf_gpu_instance_destroy / tbf_gpu_*_tables_destroy: the CLAP host frees and re-creates the
  tone generator on every tuning change (src/clap.cpp:129-160: freeToneGenerator + allocTonegen
  + initToneGenerator); 1000 rounds of that pattern must keep device memory and the instance
  array flat, both through the C-ABI and through the reference-named shim.
* TBF_RENDER_STEMS against whirlProc2's outHL / outHR / outDL / outDR (src/whirl.h:232-238).
* TBF_RENDER_OUT_CHUNKED (chunk-major host layout) against the row layout.
* two contexts in one process (one per GPU when the box has two): wide-geometry and
  bypass + checksum batches need more than 48 KB of shared memory, which is opted into per device.
"""
import ctypes as C
import os

import numpy as np
import pytest

from tunebfree_b200 import cabi, engine, tables, workloads

pytestmark = pytest.mark.gpu


def _bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def _free_bytes():
    import torch
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def test_retune_pattern_keeps_memory_flat(ref):
    """free -> alloc -> init -> render, 1000 times (src/clap.cpp:129-160)."""
    ctx = engine.Context(0)
    tt, wt = tables.InstanceSpec(seed=1).build()
    wh = ctx.whirl_tables(wt)
    ev = workloads.instance_events(3, 2, config=3)
    # a resident organ that must survive everything around it, bit for bit
    tg0 = ctx.tonegen_tables(tt)
    keep = ctx.create_instances(tg0, wh, 1)
    first_l, _ = ctx.render(1, 8, [workloads.instance_events(1, 8, config=3)], cabi.RENDER_WHIRLPROC3, ids=[keep])
    free0 = None
    for it in range(1000):
        tg = ctx.tonegen_tables(tt)                      # initToneGenerator: new wave tables
        i = ctx.create_instances(tg, wh, 1)
        if it % 50 == 0:
            ctx.render(1, 2, [ev], cabi.RENDER_WHIRLPROC3, ids=[i])
        ctx.destroy_instances(i, 1)                      # freeToneGenerator
        ctx.destroy_tonegen_tables(tg)
        live, slots, n_tg, n_wh = ctx.resource_counts()
        assert (live, n_tg, n_wh) == (1, 1, 1) and slots <= 2
        assert tg <= 1 and i <= 1                        # handles and slots are recycled
        if it == 10:
            free0 = _free_bytes()
    assert abs(_free_bytes() - free0) <= (8 << 20), "device memory grew over 1000 re-tunes"
    # the survivor still plays from where it was
    r = ref.RefInstance(seed=1)
    rl, _ = r.render(16, cabi.RENDER_WHIRLPROC3, workloads.instance_events(1, 8, config=3))
    again_l, _ = ctx.render(1, 8, [[]], cabi.RENDER_WHIRLPROC3, ids=[keep])
    assert np.array_equal(_bits(first_l[0]), _bits(rl[:8 * 128]))
    assert np.array_equal(_bits(again_l[0]), _bits(rl[8 * 128:]))
    r.close()
    # order does not matter: tables destroyed first stay alive until their last instance goes
    tg = ctx.tonegen_tables(tt)
    i = ctx.create_instances(tg, wh, 4)
    ctx.destroy_tonegen_tables(tg)
    with pytest.raises(engine.TbfError):
        ctx.create_instances(tg, wh, 1)                  # released: no new users
    ctx.render(4, 2, [ev] * 4, cabi.RENDER_WHIRLPROC3, ids=list(range(i, i + 4)))
    assert ctx.resource_counts()[2] == 2
    ctx.destroy_instances(i, 4)
    assert ctx.resource_counts() == (1, 1, 1, 1)
    with pytest.raises(engine.TbfError):
        ctx.render(1, 2, [ev], cabi.RENDER_WHIRLPROC3, ids=[i])   # destroyed instance
    with pytest.raises(engine.TbfError):
        ctx.destroy_instances(i, 1)
    tables.free_tables(tt, wt)
    ctx.close()


def test_shim_free_releases_device_memory():
    """the same pattern through the reference's own names (libtbf_shim.so)."""
    from tests.test_shim import LIB
    L = C.CDLL(LIB)
    L.allocTonegen.restype = C.c_void_p
    L.allocWhirl.restype = C.c_void_p
    buf = (C.c_float * 128)()
    w = L.allocWhirl()
    L.initWhirl(C.c_void_p(w), None, C.c_double(48000.0))
    free0 = None
    for it in range(200):
        t = L.allocTonegen()
        L.initToneGenerator(C.c_void_p(t), None, C.c_double(48000.0), None)
        if it % 40 == 0:
            L.oscKeyOn(C.c_void_p(t), 60, 60)
            L.oscGenerateFragment(C.c_void_p(t), buf, 128)
        L.freeToneGenerator(C.c_void_p(t))
        if it == 5:
            free0 = _free_bytes()
    assert L.tbf_shim_status() == 0
    assert abs(_free_bytes() - free0) <= (8 << 20)
    L.freeWhirl(C.c_void_p(w))
    L.tbf_shim_shutdown()


def test_stems_match_whirlproc2(ref, gpu_ctx):
    """TBF_RENDER_STEMS: outHL outHR outDL outDR next to the plain whirlProc mix, and the mix
    unchanged by asking for them; also a bypassed Leslie (whirl.bypass=1)."""
    rng = np.random.RandomState(11)
    n_frag = 90
    x = (rng.standard_normal(n_frag * 128) * 0.25).astype(np.float32)
    for cfg in ("", "whirl.bypass=1\n"):
        r = ref.RefInstance(seed=1, cfg=cfg)
        tg = gpu_ctx.tonegen_tables(r.tonegen_tables())
        wh = gpu_ctx.whirl_tables(r.whirl_tables())
        i = gpu_ctx.create_instances(tg, wh, 1)
        rl, rr, rst = r.whirl2(n_frag, x)
        gl, gr, gst = gpu_ctx.render(1, n_frag, [[]], 0, ids=[i], inputs=x[None, :], stems=True)
        assert np.array_equal(_bits(gl[0]), _bits(rl)) and np.array_equal(_bits(gr[0]), _bits(rr)), cfg
        for j, name in enumerate(("HL", "HR", "DL", "DR")):
            assert np.array_equal(_bits(gst[j][0]), _bits(rst[j])), (cfg, name)
        if not cfg:
            assert np.abs(rst[0]).max() > 1e-3 and np.abs(rst[2]).max() > 1e-3
        r.close()
    # full chain with stems: L / R identical to the render without them
    ev = workloads.instance_events(2, 40, config=3)
    r = ref.RefInstance(seed=1)
    tg = gpu_ctx.tonegen_tables(r.tonegen_tables())
    wh = gpu_ctx.whirl_tables(r.whirl_tables())
    a, b = gpu_ctx.create_instances(tg, wh, 1), gpu_ctx.create_instances(tg, wh, 1)
    l0, r0 = gpu_ctx.render(1, 40, [ev], cabi.RENDER_WHIRLPROC3, ids=[a])
    l1, r1, _ = gpu_ctx.render(1, 40, [ev], cabi.RENDER_WHIRLPROC3, ids=[b], stems=True)
    assert np.array_equal(_bits(l0), _bits(l1)) and np.array_equal(_bits(r0), _bits(r1))
    r.close()


def test_chunk_major_host_layout(ref, gpu_ctx):
    """TBF_RENDER_OUT_CHUNKED: chunk c of instance i at (c * n + i) * chunk_samples."""
    n, n_frag = 5, 300
    r = ref.RefInstance(seed=1)
    tg = gpu_ctx.tonegen_tables(r.tonegen_tables())
    wh = gpu_ctx.whirl_tables(r.whirl_tables())
    r.close()
    evs = [workloads.instance_events(g, n_frag, config=3) for g in range(n)]
    a = gpu_ctx.create_instances(tg, wh, n)
    b = gpu_ctx.create_instances(tg, wh, n)
    l0, r0 = gpu_ctx.render(n, n_frag, evs, cabi.RENDER_WHIRLPROC3, ids=list(range(a, a + n)))
    chunk = gpu_ctx.host_chunk()
    cs = chunk * 128
    n_chunks = -(-n_frag // chunk)
    hl, pl = gpu_ctx.host_alloc((n_chunks, n, cs))
    hr, pr = gpu_ctx.host_alloc((n_chunks, n, cs))
    gpu_ctx.prepare(n, n_frag, engine.events_to_array(evs), cabi.RENDER_WHIRLPROC3 | cabi.RENDER_OUT_CHUNKED,
                    hl.ctypes.data, hr.ctypes.data, 0, ids=list(range(b, b + n)))
    gpu_ctx.launch()
    gpu_ctx.synchronize()
    l1 = hl.transpose(1, 0, 2).reshape(n, n_chunks * cs)[:, :n_frag * 128]
    r1 = hr.transpose(1, 0, 2).reshape(n, n_chunks * cs)[:, :n_frag * 128]
    assert np.array_equal(_bits(l0), _bits(l1)) and np.array_equal(_bits(r0), _bits(r1))
    gpu_ctx.host_free(pl)
    gpu_ctx.host_free(pr)


def test_two_contexts_in_one_process(ref):
    """One context per GPU in one process (both on GPU 0 when the box has one): batches that need
    more than 48 KB of dynamic shared memory -- a wide Leslie geometry (ring 2048) and a bypassed
    Leslie with checksums -- on the SECOND context, after the first one has run them."""
    import torch
    devs = [0, 1 if torch.cuda.device_count() > 1 else 0]
    wide = "whirl.horn.radius=45\nwhirl.drum.radius=50\nwhirl.mic.distance=70\n"
    n_frag = 24
    ev = workloads.instance_events(4, n_frag, config=3, restrike=20, hold=12)
    want = {}
    for cfg in ("", wide):
        r = ref.RefInstance(seed=1, cfg=cfg)
        tt, wt = r.tonegen_tables(), r.whirl_tables()  # before rendering: they carry the initial state
        want[cfg] = (r.render(n_frag, cabi.RENDER_WHIRLPROC3, ev), tt, wt, r)
    r2 = ref.RefInstance(seed=1)
    mono = r2.render(n_frag, cabi.RENDER_WHIRL_BYPASS, ev)[0]
    r2.close()
    ctxs = [engine.Context(d) for d in devs]
    for ctx in ctxs:
        for cfg in ("", wide):
            (rl, rr), tt, wt, _ = want[cfg]
            i = ctx.create_instances(ctx.tonegen_tables(tt), ctx.whirl_tables(wt), 1)
            gl, gr = ctx.render(1, n_frag, [ev], cabi.RENDER_WHIRLPROC3, ids=[i])
            assert np.array_equal(_bits(gl[0]), _bits(rl)) and np.array_equal(_bits(gr[0]), _bits(rr)), (ctx.device, cfg)
        (_, tt, wt, _) = want[""]
        i = ctx.create_instances(ctx.tonegen_tables(tt), ctx.whirl_tables(wt), 1)
        gl, _, cs = ctx.render(1, n_frag, [ev], cabi.RENDER_WHIRL_BYPASS, ids=[i], checksums=True)
        assert np.array_equal(_bits(gl[0]), _bits(mono))
        assert int(cs[0]) == engine.checksum_host(mono, mono)
    for ctx in ctxs:
        ctx.close()
    for cfg in want:
        want[cfg][3].close()


def test_event_list_checks_and_the_threaded_event_index(gpu_ctx):
    """tbf_gpu_batch_prepare rejects event lists that are not sorted by (slot, fragment) or name a slot /
    fragment outside the batch, and its per-slot index is the same whether one thread builds it or
    several do (lists of 2^18 events and more are cut into ranges): one batch of 330 000 events against
    the same organs rendered as two batches of half the organs."""
    tt, wt = tables.InstanceSpec(seed=2).build()
    tg, wh = gpu_ctx.tonegen_tables(tt), gpu_ctx.whirl_tables(wt)
    tables.free_tables(tt, wt)
    n, n_frag = 44, 300
    a = gpu_ctx.create_instances(tg, wh, n)
    on, off = cabi.EV_KEY_ON, cabi.EV_KEY_OFF
    for bad in ([[0, 5, on, 30, 0], [0, 4, off, 30, 0]],            # fragments go back inside a slot
                [[0, 0, on, 30, 0], [0, n_frag, off, 30, 0]],       # fragment past the end
                [[1, 0, on, 30, 0], [0, 1, off, 30, 0]],            # slots go back
                [[0, 0, on, 30, 0], [2, 1, off, 30, 0]]):           # slot past the batch
        with pytest.raises(engine.TbfError):
            gpu_ctx.render(2, n_frag, engine.events_from_numpy(np.array(bad, dtype=np.int32)),
                           cabi.RENDER_WHIRLPROC3, ids=[a, a + 1])
    good = np.array([[0, 0, on, 30, 0], [0, 5, off, 30, 0], [1, 5, on, 31, 0]], dtype=np.int32)
    gpu_ctx.render(2, n_frag, engine.events_from_numpy(good), cabi.RENDER_WHIRLPROC3, ids=[a, a + 1])
    gpu_ctx.destroy_instances(a, n)

    rng = np.random.RandomState(5)
    per = 7500                                   # events per organ: 44 x 7500 = 330 000 > 2^18
    rows = []
    for g in range(n):
        if g in (0, 17):                         # organs without any event, first and in the middle
            continue
        fr = np.sort(rng.randint(0, n_frag, per))
        ev = np.empty((per, 5), dtype=np.int32)
        ev[:, 0], ev[:, 1] = g, fr
        ev[:, 2] = cabi.EV_DRAWBAR
        ev[:, 3] = rng.randint(0, 27, per)
        ev[:, 4] = rng.randint(0, 9, per)
        ev[:40, 1], ev[:40, 2] = 0, cabi.EV_KEY_ON
        ev[:40, 3] = np.concatenate([rng.choice(61, 20, replace=False), 64 + rng.choice(61, 20, replace=False)])
        ev[:40, 4] = 0
        rows.append(ev)
    ev_all = np.concatenate(rows)
    assert ev_all.shape[0] >= (1 << 18)
    a = gpu_ctx.create_instances(tg, wh, n)
    l0, r0 = gpu_ctx.render(n, n_frag, engine.events_from_numpy(ev_all), cabi.RENDER_WHIRLPROC3,
                            ids=list(range(a, a + n)))
    gpu_ctx.destroy_instances(a, n)
    assert np.abs(l0[1]).max() > 0.01 and np.abs(l0[0]).max() < 1e-9   # organ 0 never plays a key
    b = gpu_ctx.create_instances(tg, wh, n)
    h = n // 2
    for lo, hi in ((0, h), (h, n)):
        sub = ev_all[(ev_all[:, 0] >= lo) & (ev_all[:, 0] < hi)].copy()
        sub[:, 0] -= lo
        assert sub.shape[0] < (1 << 18)
        l1, r1 = gpu_ctx.render(hi - lo, n_frag, engine.events_from_numpy(sub), cabi.RENDER_WHIRLPROC3,
                                ids=list(range(b + lo, b + hi)))
        assert np.array_equal(_bits(l1), _bits(l0[lo:hi])) and np.array_equal(_bits(r1), _bits(r0[lo:hi]))
    gpu_ctx.destroy_instances(b, n)
