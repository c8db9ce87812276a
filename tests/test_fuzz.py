"""Seeded random performances against the oracle, bit for bit.

Each seed plays a few hundred fragments of random key presses and releases on all three manuals,
drawbar, percussion, vibrato-routing, swell and rotor-speed changes at random fragments -- the
control surface of the hot path (src/tonegen.cpp setters, src/whirl.cpp useRevOption) -- through
the reference engine and through the CUDA path, with the default kernels and with stage T's other
sample movers.  The workloads in the other parity tests are hand-made; these are not, so the quiet /
full-pass transitions, the bus masks and lists, the record store across launches and the ordered
scatter see combinations nobody wrote down.
"""
import numpy as np
import pytest

from tunebfree_b200 import cabi

pytestmark = pytest.mark.gpu


def _bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def _performance(seed, n_frag):
    rng = np.random.RandomState(seed)
    ev = [(0, cabi.EV_DRAWBAR, b, int(rng.randint(0, 9))) for b in range(20)]
    ev += [(0, cabi.EV_REV_OPTION, int(rng.randint(0, 9)), 0)]
    held = set()
    f = 0
    while f < n_frag:
        kind = rng.randint(0, 100)
        if kind < 45:                                    # press 1..6 keys, any manual
            for _ in range(rng.randint(1, 7)):
                base = (0, 64, 128)[rng.randint(0, 3)]
                k = base + int(rng.randint(0, 25 if base == 128 else 61))
                if k not in held:
                    held.add(k)
                    ev.append((f, cabi.EV_KEY_ON, k, 0))
        elif kind < 70 and held:                         # release some
            for k in list(held)[: rng.randint(1, 5)]:
                held.discard(k)
                ev.append((f, cabi.EV_KEY_OFF, k, 0))
        elif kind < 78:
            ev.append((f, cabi.EV_DRAWBAR, int(rng.randint(0, 27)), int(rng.randint(0, 9))))
        elif kind < 83:
            ev.append((f, cabi.EV_PERC_ENABLE, int(rng.randint(0, 2)), 0))
        elif kind < 86:
            ev.append((f, (cabi.EV_PERC_SOFT, cabi.EV_PERC_FAST, cabi.EV_PERC_FIRST)[rng.randint(0, 3)],
                       int(rng.randint(0, 2)), 0))
        elif kind < 91:
            ev.append((f, (cabi.EV_VIB_UPPER, cabi.EV_VIB_LOWER)[rng.randint(0, 2)], int(rng.randint(0, 2)), 0))
        elif kind < 94:
            ev.append((f, cabi.EV_VIB_KNOB, int(rng.randint(0, 6)), 0))
        elif kind < 97:
            ev.append((f, cabi.EV_SWELL, int(rng.randint(20, 128)), 0))
        else:
            ev.append((f, cabi.EV_REV_OPTION, int(rng.randint(0, 9)), 0))
        f += int(rng.choice([1, 1, 2, 3, 5, 9, 17, 40]))  # bursts and long quiet stretches
    ev.sort(key=lambda e: e[0])
    return ev


@pytest.mark.parametrize("seed", list(range(11, 27)))
def test_random_performance_is_bit_exact(ref, gpu_ctx, seed):
    n_frag = 420
    ev = _performance(seed, n_frag)
    flags = cabi.RENDER_WHIRLPROC3 if seed % 3 else cabi.RENDER_WHIRL_BYPASS
    r = ref.RefInstance(seed=1 + seed % 4)
    tg = gpu_ctx.tonegen_tables(r.tonegen_tables())
    wh = gpu_ctx.whirl_tables(r.whirl_tables())
    rl, rr = r.render(n_frag, flags, ev)
    rs = r.export_state()
    r.close()
    for extra in (0, cabi.RENDER_DEBUG_T_TMA, cabi.RENDER_DEBUG_T_CPASYNC | cabi.RENDER_DEBUG_NO_TMA_TILES,
                  cabi.RENDER_DEBUG_T_DIRECT | cabi.RENDER_DEBUG_BUS_ALL):
        i = gpu_ctx.create_instances(tg, wh, 1)
        # in three calls: the record store and every other state crosses launches
        parts = []
        for f0, f1 in ((0, 150), (150, 151), (151, n_frag)):
            sub = [(f - f0, t, a, b) for (f, t, a, b) in ev if f0 <= f < f1]
            parts.append(gpu_ctx.render(1, f1 - f0, [sub], flags | extra, ids=[i]))
        gl = np.concatenate([p[0][0] for p in parts])
        gr = np.concatenate([p[1][0] for p in parts])
        bad = np.nonzero(_bits(gl) != _bits(rl))[0]
        assert bad.size == 0, "seed %d variant %#x: L differs first at sample %d (fragment %d)" % (
            seed, extra, bad[0], bad[0] // 128)
        assert np.array_equal(_bits(gr), _bits(rr)), (seed, extra)
        gs = gpu_ctx.export_state(i)
        assert list(gs.osc_pos) == list(rs.osc_pos) and list(gs.ref_count) == list(rs.ref_count)
        assert gs.active_osc_l_end == rs.active_osc_l_end and gs.error_flags == 0
        gpu_ctx.destroy_instances(i, 1)


def test_forty_random_organs_in_one_batch(ref, gpu_ctx):
    """40 different performances in ONE batch (a full 32-organ tile and a ragged one of 8), two table
    sets, rendered in two calls; every organ against its own oracle run."""
    n_frag, n = 160, 40
    r0 = ref.RefInstance(seed=1)
    r1 = ref.RefInstance(seed=2)
    tgs = [gpu_ctx.tonegen_tables(r0.tonegen_tables()), gpu_ctx.tonegen_tables(r1.tonegen_tables())]
    wh = gpu_ctx.whirl_tables(r0.whirl_tables())
    r0.close()
    r1.close()
    evs = [_performance(100 + g, n_frag) for g in range(n)]
    ids = [gpu_ctx.create_instances(tgs[g % 2], wh, 1) for g in range(n)]
    parts = []
    for f0, f1 in ((0, 70), (70, n_frag)):
        sub = [[(f - f0, t, a, b) for (f, t, a, b) in ev if f0 <= f < f1] for ev in evs]
        parts.append(gpu_ctx.render(n, f1 - f0, sub, cabi.RENDER_WHIRLPROC3, ids=ids))
    gl = np.concatenate([p[0] for p in parts], axis=1)
    gr = np.concatenate([p[1] for p in parts], axis=1)
    for g in range(n):
        r = ref.RefInstance(seed=1 + g % 2)
        rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, evs[g])
        r.close()
        assert np.array_equal(_bits(gl[g]), _bits(rl)) and np.array_equal(_bits(gr[g]), _bits(rr)), g
    for i in ids:
        gpu_ctx.destroy_instances(i, 1)
