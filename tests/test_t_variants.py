"""Stage T's variants all produce the reference's bits.

How the wheels' samples reach the sums is a launch-time choice (TBF_RENDER_DEBUG_T_*: TMA bulk
copies, per-lane cp.async, direct register loads), and with direct loads a quiet fragment is summed
bus by bus over the records whose gain on that bus is not zero (src/tonegen.cpp:3640-3680 multiplies
by the zeros too).  Every combination is compared bit for bit with the oracle on workloads that
exercise each case of the bus-by-bus path: a swell bus nobody feeds (all keys on a manual routed to
the scanner), one fed by some records (lower manual without vibrato next to the upper one with),
no scanner routing at all, percussion on and off, silence between chords, drawbar changes mid-note.
Stages R and D move their organ x sample tiles by 2-D TMA boxes or (TBF_RENDER_DEBUG_NO_TMA_TILES)
by cp.async and store loops: both are in the combinations.
"""
import numpy as np
import pytest

from tunebfree_b200 import cabi, workloads

pytestmark = pytest.mark.gpu

MOVERS = {"tma": 0x1000, "cpasync": 0x2000, "direct": 0x3000}
BUS_ALL, NO_BUS, NO_PAIRS, ALL_BUSES, NO_TILES = 0x400, 0x800, 0x200, 0x4000, 0x8000


def _bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def _workload(kind, n_frag):
    if kind == "config3":
        return workloads.instance_events(3, n_frag, config=3, restrike=60, hold=45), cabi.RENDER_WHIRLPROC3
    if kind == "config2":
        return workloads.instance_events(7, n_frag, config=2, restrike=60, hold=45), cabi.RENDER_WHIRL_BYPASS
    ev = [(0, cabi.EV_DRAWBAR, b, 8) for b in range(9)]            # upper 888888888
    ev += [(0, cabi.EV_DRAWBAR, 9 + b, 6) for b in range(9)]       # lower
    ev += [(0, cabi.EV_DRAWBAR, 18, 8), (0, cabi.EV_DRAWBAR, 19, 5)]  # pedal
    if kind == "upper_only_vib":
        # every key on the upper manual, routed to the scanner: no record has a swell gain
        ev += [(0, cabi.EV_VIB_UPPER, 1, 0), (0, cabi.EV_VIB_KNOB, 5, 0), (0, cabi.EV_PERC_ENABLE, 1, 0)]
        keys = [24, 31, 36, 40, 43, 48, 52]
    elif kind == "mixed_manuals":
        # upper with vibrato, lower and pedal without: wheels shared by the manuals feed both buses
        ev += [(0, cabi.EV_VIB_UPPER, 1, 0), (0, cabi.EV_VIB_KNOB, 3, 0), (0, cabi.EV_PERC_ENABLE, 1, 0)]
        keys = [24, 36, 43, 64 + 12, 64 + 19, 64 + 24, 64 + 36, 128 + 5]
    elif kind == "no_perc_no_vib":
        ev += [(0, cabi.EV_PERC_ENABLE, 0, 0)]
        keys = [30, 37, 42, 64 + 30, 128 + 2]
    else:
        raise ValueError(kind)
    f = 0
    while f < n_frag:
        ev += [(f, cabi.EV_KEY_ON, k, 0) for k in keys]
        if f + 50 < n_frag:
            ev += [(f + 50, cabi.EV_KEY_OFF, k, 0) for k in keys]
        if f + 21 < n_frag:
            ev.append((f + 21, cabi.EV_DRAWBAR, 2, (f // 7) % 9))  # drawbar change while held
        if kind == "mixed_manuals" and f + 33 < n_frag:
            ev.append((f + 33, cabi.EV_VIB_LOWER, (f // 70) % 2, 0))  # routing change while held
        f += 70
    ev.sort(key=lambda e: e[0])
    return ev, cabi.RENDER_WHIRLPROC3 if kind != "no_perc_no_vib" else cabi.RENDER_WHIRL_BYPASS


KINDS = ["config3", "config2", "upper_only_vib", "mixed_manuals", "no_perc_no_vib"]


@pytest.mark.parametrize("kind", KINDS)
def test_every_mover_and_bus_mode_is_bit_exact(ref, gpu_ctx, kind):
    n_frag = 300
    ev, flags = _workload(kind, n_frag)
    r = ref.RefInstance(seed=2)
    tg = gpu_ctx.tonegen_tables(r.tonegen_tables())
    wh = gpu_ctx.whirl_tables(r.whirl_tables())
    rl, rr = r.render(n_frag, flags, ev)
    rs = r.export_state()
    r.close()
    assert np.abs(rl).max() > 0.01
    combos = [(m, 0) for m in MOVERS] + [("direct", BUS_ALL), ("direct", NO_BUS), ("direct", NO_PAIRS),
                                         ("direct", NO_PAIRS | BUS_ALL), ("direct", ALL_BUSES), ("direct", ALL_BUSES | NO_BUS),
                                         ("tma", NO_PAIRS), ("direct", NO_TILES), ("tma", NO_TILES)]
    for mover, extra in combos:
        i = gpu_ctx.create_instances(tg, wh, 1)
        gl, gr = gpu_ctx.render(1, n_frag, [ev], flags | MOVERS[mover] | extra, ids=[i])
        bad = np.nonzero(_bits(gl[0]) != _bits(rl))[0]
        assert bad.size == 0, "%s / %s+%#x: L differs first at sample %d (fragment %d)" % (
            kind, mover, extra, bad[0], bad[0] // 128)
        assert np.array_equal(_bits(gr[0]), _bits(rr)), (kind, mover, extra)
        gs = gpu_ctx.export_state(i)
        assert list(gs.osc_pos) == list(rs.osc_pos) and gs.error_flags == 0
        assert np.array_equal(np.ctypeslib.as_array(gs.vib_buffer).view(np.uint32),
                              np.ctypeslib.as_array(rs.vib_buffer).view(np.uint32)), (kind, mover, extra)
        gpu_ctx.destroy_instances(i, 1)


def test_many_organs_direct_against_tma(gpu_ctx):
    """A full-size batch (more CTAs than fit at once): the default mover against the TMA one,
    device checksums of every organ."""
    from tunebfree_b200 import tables
    tt, wt = tables.InstanceSpec(seed=3).build()
    tg, wh = gpu_ctx.tonegen_tables(tt), gpu_ctx.whirl_tables(wt)
    tables.free_tables(tt, wt)
    n, n_frag = 4200, 96
    evs = [workloads.instance_events(g, n_frag, config=3, restrike=40, hold=30) for g in range(n)]
    sums = []
    for mover in (0, MOVERS["tma"] | NO_TILES, MOVERS["direct"] | BUS_ALL):
        first = gpu_ctx.create_instances(tg, wh, n)
        _, _, cs = gpu_ctx.render(n, n_frag, evs, cabi.RENDER_WHIRLPROC3 | cabi.RENDER_NO_OUTPUT | mover,
                                  ids=list(range(first, first + n)), checksums=True)
        sums.append(np.array(cs, dtype=np.uint64))
        gpu_ctx.destroy_instances(first, n)
    assert np.array_equal(sums[0], sums[1]) and np.array_equal(sums[0], sums[2])
