"""GPU parity tests: the CUDA path, called through the C-ABI, against the oracle
(the reference engine compiled in place) on the same seeded inputs.

Bar: BIT-EXACT float32 output (north_star's tolerance is 1e-5 of full scale; the
Leslie's DF-II biquads amplify 1-ulp differences past that, SURVEY 6.2, so the
design target is 0 and these tests assert array equality), and bit-exact integer
state (wave positions, active list, reference counts, delay-line indices).
"""
import ctypes as C

import numpy as np
import pytest

from tunebfree_b200 import cabi, workloads

pytestmark = pytest.mark.gpu


def _make(ref, ctx, n=1, seed=1, cfg="", freqs=None, rate=48000.0):
    r = ref.RefInstance(freqs=freqs, seed=seed, cfg=cfg, rate=rate)
    tg = ctx.tonegen_tables(r.tonegen_tables())
    wh = ctx.whirl_tables(r.whirl_tables())
    first = ctx.create_instances(tg, wh, n)
    return r, first


def _compare_state(gs, rs, whirl=True):
    n = rs.active_osc_l_end
    assert gs.active_osc_l_end == n
    assert list(gs.active_osc_list)[:n] == list(rs.active_osc_list)[:n]
    assert list(gs.osc_pos) == list(rs.osc_pos)
    assert list(gs.osc_acl_pos)[1:] == list(rs.osc_acl_pos)[1:]
    assert list(gs.ref_count) == list(rs.ref_count)
    assert (gs.old_routing, gs.new_routing, gs.upper_key_count, gs.key_down_count) == \
           (rs.old_routing, rs.new_routing, rs.upper_key_count, rs.key_down_count)
    f32 = lambda x: np.float32(x).view(np.uint32)
    assert f32(gs.perc_env_gain) == f32(rs.perc_env_gain)
    assert f32(gs.key_comp_level) == f32(rs.key_comp_level)
    assert f32(gs.pz) == f32(rs.pz)
    assert (gs.vib_stator, gs.vib_out_pos) == (rs.vib_stator, rs.vib_out_pos)
    assert np.array_equal(np.ctypeslib.as_array(gs.vib_buffer).view(np.uint32),
                          np.ctypeslib.as_array(rs.vib_buffer).view(np.uint32))
    if whirl:
        assert gs.whirl_outpos == rs.whirl_outpos
        assert (gs.horn_angle, gs.drum_angle, gs.horn_incr, gs.drum_incr) == \
               (rs.horn_angle, rs.drum_angle, rs.horn_incr, rs.drum_incr)
        assert (gs.horn_acdc, gs.drum_acdc) == (rs.horn_acdc, rs.drum_acdc)
        assert tuple(gs.brake_pos) == tuple(rs.brake_pos)
        assert tuple(gs.ramp_one_minus_l) == tuple(rs.ramp_one_minus_l)
        for name in ("whirl_buf", "iir_z", "horn_iir", "filt_z", "adx"):
            a = np.ctypeslib.as_array(getattr(gs, name)).view(np.uint32)
            b = np.ctypeslib.as_array(getattr(rs, name)).view(np.uint32)
            assert np.array_equal(a, b), name
        assert list(gs.adi) == list(rs.adi)
    assert gs.error_flags == 0


def _bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def _check(ref, ctx, events, n_frag, flags, seed=1, cfg="", freqs=None, whirl_state=True, rate=48000.0):
    r, first = _make(ref, ctx, 1, seed, cfg, freqs, rate)
    rl, rr = r.render(n_frag, flags, events)
    gl, gr, cs = ctx.render(1, n_frag, [events], flags, ids=[first], checksums=True)
    bad = np.nonzero(_bits(gl[0]) != _bits(rl))[0]
    assert bad.size == 0, "L differs first at sample %d (fragment %d): gpu %r ref %r, max abs %g" % (
        bad[0], bad[0] // 128, gl[0][bad[0]], rl[bad[0]], np.abs(gl[0] - rl).max())
    assert np.array_equal(_bits(gr[0]), _bits(rr))
    from tunebfree_b200.engine import checksum_host
    assert int(cs[0]) == checksum_host(gl[0], gr[0])
    _compare_state(ctx.export_state(first), r.export_state(), whirl_state)
    peak = float(np.abs(rl).max())
    r.close()
    return peak


def test_tonegen_only_chord_click_percussion(ref, gpu_ctx):
    """BASELINE config 2 shape: full drawbars, 10-note chord re-struck, percussion +
    key click, whirl bypassed."""
    ev = workloads.instance_events(7, 600, config=2)
    peak = _check(ref, gpu_ctx, ev, 600, cabi.RENDER_WHIRL_BYPASS, whirl_state=False)
    assert peak > 0.05


def test_full_chain_config3(ref, gpu_ctx):
    """BASELINE config 3 shape: tonegen + vibrato C3 + Leslie fast through whirlProc3."""
    ev = workloads.instance_events(3, 800, config=3)
    peak = _check(ref, gpu_ctx, ev, 800, cabi.RENDER_WHIRLPROC3)
    assert peak > 0.05


@pytest.mark.parametrize("rate", [44100.0, 22050.0, 96000.0])
def test_other_sample_rates(ref, gpu_ctx, rate):
    """The hosts run the engine at the session's rate (SampleRateD): wave lengths, envelopes,
    scanner increment, Leslie geometry and filters all scale with it."""
    ev = workloads.instance_events(11, 500, config=3, restrike=120, hold=80)
    ev += [(260, cabi.EV_REV_OPTION, 4, 0), (380, cabi.EV_REV_OPTION, 8, 0)]
    ev.sort(key=lambda e: e[0])
    peak = _check(ref, gpu_ctx, ev, 500, cabi.RENDER_WHIRLPROC3, rate=rate)
    assert peak > 0.05


def test_whirlproc_plain_and_speed_changes(ref, gpu_ctx):
    """whirlProc (no mic mix); all nine useRevOption speeds incl. accel/decel ramps."""
    ev = workloads.setup_events(3)
    ev += [(0, cabi.EV_KEY_ON, k, 0) for k in (48, 55, 60, 64)]
    for i, n in enumerate([4, 8, 0, 2, 6, 1, 7, 3, 5, 8]):
        ev.append((40 + 90 * i, cabi.EV_REV_OPTION, n, 0))
    ev += [(700, cabi.EV_REV_SELECT, 1, 0), (820, cabi.EV_REV_SELECT, 2, 0)]
    _check(ref, gpu_ctx, ev, 1000, 0)


def test_leslie_controllers_brakes_ramps_horn_filters(ref, gpu_ctx):
    """TBF_EV_WHIRL_BRAKE / _RAMP / TBF_EV_HORN_FILTER (src/whirl.cpp:703-917): brake positions
    while the rotors run down to a stop, every ramp constant at both ends of its range, horn
    biquads swapped mid-note; per-instance, so a second organ of the same table set keeps the
    table's values."""
    from tunebfree_b200 import tables
    ev = workloads.setup_events(3)
    ev += [(0, cabi.EV_KEY_ON, k, 0) for k in (40, 52, 60, 67, 76)]
    ev += [(2, cabi.EV_WHIRL_BRAKE, 0, 64), (2, cabi.EV_WHIRL_BRAKE, 1, 127), (3, cabi.EV_WHIRL_RAMP, 1, 0),
           (3, cabi.EV_WHIRL_RAMP, 3, 5), (10, cabi.EV_REV_OPTION, 0, 0),       # both rotors to a stop
           (330, cabi.EV_WHIRL_RAMP, 0, 127), (330, cabi.EV_WHIRL_RAMP, 2, 0), (331, cabi.EV_REV_OPTION, 8, 0),
           (480, cabi.EV_WHIRL_BRAKE, 0, 1), (480, cabi.EV_WHIRL_BRAKE, 1, 0), (481, cabi.EV_REV_SELECT, 1, 0),
           (600, cabi.EV_WHIRL_RAMP, 1, 127), (600, cabi.EV_WHIRL_RAMP, 3, 127), (601, cabi.EV_REV_SELECT, 0, 0)]
    for frag, filt, (typ, hz, q, g) in [(100, 0, (1, 1200.0, 0.8, 0.0)), (200, 1, (2, 700.0, 3.0, 0.0)),
                                        (400, 0, (8, 5000.0, 1.0, -20.0)), (640, 1, (3, 2500.0, 5.5, 12.0))]:
        c5 = tables.iir_coefficients(typ, hz, q, g, 48000.0)
        ev += [(frag, cabi.EV_HORN_FILTER, 8 * filt + k, int(np.float32(c5[k]).view(np.int32))) for k in range(5)]
    ev.sort(key=lambda e: e[0])
    _check(ref, gpu_ctx, ev, 760, cabi.RENDER_WHIRLPROC3)
    # two organs of one table set: only the first gets the controllers
    r, first = _make(ref, gpu_ctx, 2)
    plain = workloads.setup_events(3) + [(0, cabi.EV_KEY_ON, 60, 0), (5, cabi.EV_REV_OPTION, 0, 0)]
    gl, gr = gpu_ctx.render(2, 200, [[e for e in ev if e[0] < 200], plain], cabi.RENDER_WHIRLPROC3,
                            ids=[first, first + 1])
    rl, rr = r.render(200, cabi.RENDER_WHIRLPROC3, plain)
    r.close()
    assert np.array_equal(_bits(gl[1]), _bits(rl)) and np.array_equal(_bits(gr[1]), _bits(rr))
    assert not np.array_equal(_bits(gl[0]), _bits(gl[1]))
    gpu_ctx.render(1, 4, [[(0, cabi.EV_WHIRL_RAMP, 4, 10)]], 0, ids=[first + 1])
    assert gpu_ctx.export_state(first + 1).error_flags & cabi.IFLAG_BAD_EVENT
    assert not gpu_ctx.export_state(first).error_flags & cabi.IFLAG_BAD_EVENT


def test_setters_midstream(ref, gpu_ctx):
    """Drawbar, percussion, vibrato routing / knob and swell changes between fragments,
    keys on all three manuals, re-striking a held key, silence with the scanner on."""
    E = cabi
    ev = [(0, E.EV_KEY_ON, 60, 0), (0, E.EV_KEY_ON, 128 + 48, 0), (0, E.EV_KEY_ON, 256 + 36, 0),
          (5, E.EV_DRAWBAR, 3, 8), (5, E.EV_DRAWBAR, 8, 5), (9, E.EV_PERC_ENABLE, 1, 0),
          (12, E.EV_KEY_ON, 64, 0), (12, E.EV_KEY_ON, 60, 0),  # 60 re-struck while held
          (20, E.EV_VIB_UPPER, 1, 0), (25, E.EV_VIB_KNOB, 1, 0), (30, E.EV_VIB_LOWER, 1, 0),
          (33, E.EV_PERC_SOFT, 1, 0), (35, E.EV_PERC_FAST, 0, 0), (37, E.EV_PERC_FIRST, 1, 0),
          (40, E.EV_DRAWBAR, 8, 8), (44, E.EV_SWELL, 90, 0), (50, E.EV_VIB_KNOB, 4, 0),
          (55, E.EV_KEY_OFF, 60, 0), (55, E.EV_KEY_OFF, 64, 0), (60, E.EV_KEY_ON, 72, 0),
          (70, E.EV_PERC_ENABLE, 0, 0), (75, E.EV_VIB_UPPER, 0, 0), (80, E.EV_VIB_KNOB, 3, 0),
          (90, E.EV_KEY_OFF, 72, 0), (90, E.EV_KEY_OFF, 128 + 48, 0), (90, E.EV_KEY_OFF, 256 + 36, 0),
          (100, E.EV_DRAWBAR, 9, 8), (100, E.EV_DRAWBAR, 20, 8), (130, E.EV_KEY_ON, 40, 0),
          (131, E.EV_KEY_OFF, 40, 0), (131, E.EV_KEY_ON, 40, 0), (160, E.EV_KEY_OFF, 40, 0),
          (0, E.EV_REV_OPTION, 4, 0)]
    _check(ref, gpu_ctx, ev, 260, cabi.RENDER_WHIRLPROC3)


def test_ragged_batch_silent_and_playing(ref, gpu_ctx):
    """Ragged input: organs without any event (silence still runs through the stateful scanner
    and the Leslie, src/tonegen.cpp:3600-3619), organs with control events only, organs whose
    only events sit on the last fragment, next to playing ones -- in one batch of 70 slots
    (three warps of lane = organ stages, the last one partly filled), 1 and 37 fragments."""
    r, first = _make(ref, gpu_ctx, 70)
    r.close()
    scripts = [[], workloads.setup_events(3), [(36, cabi.EV_KEY_ON, 60, 0)],
               workloads.instance_events(2, 37, config=3, restrike=12, hold=7)]
    evs = [scripts[i % 4] for i in range(70)]
    refs = []
    for sc in scripts:
        rr_ = _make(ref, gpu_ctx, 1)[0]
        a = rr_.render(1, cabi.RENDER_WHIRLPROC3, [e for e in sc if e[0] < 1])
        b = rr_.render(37, cabi.RENDER_WHIRLPROC3, [(e[0], e[1], e[2], e[3]) for e in sc])
        refs.append((a, b))
        rr_.close()
    ids = list(range(first, first + 70))
    g1 = gpu_ctx.render(70, 1, [[e for e in sc if e[0] < 1] for sc in evs], cabi.RENDER_WHIRLPROC3, ids=ids)
    g2 = gpu_ctx.render(70, 37, evs, cabi.RENDER_WHIRLPROC3, ids=ids)
    for i in range(70):
        a, b = refs[i % 4]
        for ch in (0, 1):
            assert np.array_equal(_bits(g1[ch][i]), _bits(a[ch])), (i, ch)
            assert np.array_equal(_bits(g2[ch][i]), _bits(b[ch])), (i, ch)


def test_all_61_keys_and_release(ref, gpu_ctx):
    """Many active wheels (169 wheels / 221 core instructions in the reference: close
    to its 256-instruction corePgm), then full release down to an empty active list."""
    ev = workloads.setup_events(3)
    ev += [(2, cabi.EV_KEY_ON, k, 0) for k in range(36, 97)]
    ev += [(60, cabi.EV_KEY_OFF, k, 0) for k in range(36, 97)]
    _check(ref, gpu_ctx, ev, 120, cabi.RENDER_WHIRLPROC3)


def test_long_run_wave_wraps(ref, gpu_ctx):
    """20 s held chord: every wavetable wraps many times (incl. the pos == length case)."""
    ev = workloads.setup_events(3) + [(0, cabi.EV_KEY_ON, k, 0) for k in (36, 41, 52, 60, 67, 77, 84, 96)]
    _check(ref, gpu_ctx, ev, 7500, cabi.RENDER_WHIRLPROC3)


def test_batch_split_equals_one_shot(ref, gpu_ctx):
    """Rendering in several tbf_gpu_render_batch calls carries the state exactly."""
    n_frag = 300
    ev = workloads.instance_events(11, n_frag, config=3)
    r, first = _make(ref, gpu_ctx, 1)
    rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, ev)
    outs = []
    for lo, hi in ((0, 1), (1, 100), (100, 101), (101, 300)):
        sub = [(f - lo, t, a, b) for (f, t, a, b) in ev if lo <= f < hi]
        l, rgt = gpu_ctx.render(1, hi - lo, [sub], cabi.RENDER_WHIRLPROC3, ids=[first])
        outs.append((l[0], rgt[0]))
    gl = np.concatenate([o[0] for o in outs])
    gr = np.concatenate([o[1] for o in outs])
    assert np.array_equal(_bits(gl), _bits(rl)) and np.array_equal(_bits(gr), _bits(rr))
    _compare_state(gpu_ctx.export_state(first), r.export_state())


def test_many_instances_independent(ref, gpu_ctx):
    """64 instances with different chords in one launch == 64 oracle renders."""
    n, n_frag = 64, 200
    r0, first = _make(ref, gpu_ctx, n)
    evs = [workloads.instance_events(i, n_frag, config=3) for i in range(n)]
    gl, gr = gpu_ctx.render(n, n_frag, evs, cabi.RENDER_WHIRLPROC3, ids=list(range(first, first + n)))
    for i in (0, 1, 17, 63):
        r = ref.RefInstance(seed=1)
        rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, evs[i])
        assert np.array_equal(_bits(gl[i]), _bits(rl)), i
        assert np.array_equal(_bits(gr[i]), _bits(rr)), i
        r.close()
    # reset puts every instance back to the post-init state
    gpu_ctx.reset(first, n)
    gl2, gr2 = gpu_ctx.render(n, n_frag, evs, cabi.RENDER_WHIRLPROC3, ids=list(range(first, first + n)))
    assert np.array_equal(_bits(gl), _bits(gl2)) and np.array_equal(_bits(gr), _bits(gr2))


def test_cfg_variants(ref, gpu_ctx):
    """osc-*/whirl-* style cfg variants resolved by the reference's own parser:
    sawtooth harmonics, no crosstalk, whirl-sharp's dead horn filter, brake position,
    mic width + angle, whirl bypass."""
    cfgs = [
        "osc.harmonic.1=1.0\nosc.harmonic.2=0.5\nosc.harmonic.3=0.333\nosc.harmonic.4=0.25\n",
        "osc.compartment-crosstalk=0\nosc.transformer-crosstalk=0\nosc.terminalstrip-crosstalk=0\nosc.wiring-crosstalk=0\n",
        "whirl.horn.filter.b.gain=-60\nwhirl.horn.filter.a.hz=3000\n",
        "whirl.horn.brakepos=0.25\nwhirl.drum.brakepos=0.5\n",
        "whirl.horn.mic.angle=90\nwhirl.drum.width=0.4\nwhirl.horn.width=-0.3\n",
        "whirl.bypass=1\n",
        "osc.attack.model=cosine\nosc.release.model=click\nscanner.hz=6.1\nscanner.modulation.v3=11\n",
    ]
    for cfg in cfgs:
        ev = workloads.setup_events(3) + [(1, cabi.EV_KEY_ON, k, 0) for k in (43, 60, 64, 79)]
        ev += [(150, cabi.EV_REV_OPTION, 0, 0), (320, cabi.EV_KEY_OFF, 60, 0)]
        _check(ref, gpu_ctx, ev, 700, cabi.RENDER_WHIRLPROC3, seed=5, cfg=cfg)
        _check(ref, gpu_ctx, [e for e in ev if e[0] < 100], 100, 0, seed=5, cfg=cfg)


def test_default_geometry_uses_the_parallel_gather(ref, gpu_ctx):
    """The stock Leslie geometry must run the ordered-gather delay lines on a 512-slot
    shared-memory ring (a silent fall-back to the serial lanes would hide a perf bug)."""
    r = ref.RefInstance(seed=1)
    wh = gpu_ctx.whirl_tables(r.whirl_tables())
    gather_ok, ring, window = gpu_ctx.whirl_tables_info(wh)
    assert (gather_ok, ring) == (1, 512) and 128 < window <= 640
    r.close()


def test_wide_geometry_and_fast_rotors_fall_back_exactly(ref, gpu_ctx):
    """Geometries whose tap delay ranges overlap use the serial delay-line path, and a
    rotor spun far beyond any Leslie exercises the per-fragment monotonicity check:
    both must still match the reference bit for bit."""
    ev = workloads.setup_events(3) + [(1, cabi.EV_KEY_ON, k, 0) for k in (43, 60, 64, 79)]
    cfg = "whirl.horn.radius=45\nwhirl.drum.radius=50\nwhirl.mic.distance=70\n"
    r = ref.RefInstance(seed=2, cfg=cfg)
    wh = gpu_ctx.whirl_tables(r.whirl_tables())
    assert gpu_ctx.whirl_tables_info(wh)[0] == 0
    r.close()
    _check(ref, gpu_ctx, ev, 300, cabi.RENDER_WHIRLPROC3, seed=2, cfg=cfg)
    cfg = "whirl.horn.fastrpm=90000\nwhirl.drum.fastrpm=70000\nwhirl.horn.acceleration=0.01\n"
    r, first = _make(ref, gpu_ctx, 1, 2, cfg)
    rl, rr = r.render(300, cabi.RENDER_WHIRLPROC3, ev)
    gl, gr = gpu_ctx.render(1, 300, [ev], cabi.RENDER_WHIRLPROC3, ids=[first])
    assert np.array_equal(_bits(gl[0]), _bits(rl)) and np.array_equal(_bits(gr[0]), _bits(rr))
    assert gpu_ctx.export_state(first).error_flags & cabi.IFLAG_GATHER_FALLBACK
    r.close()


def test_error_paths(gpu_ctx):
    from tunebfree_b200.engine import TbfError
    with pytest.raises(TbfError):
        gpu_ctx.render(1, 4, [[]], 0, ids=[10 ** 6])
    with pytest.raises(TbfError):
        gpu_ctx.create_instances(999, 999, 1)
    with pytest.raises(TbfError):   # Leslie-only and Leslie-bypass exclude each other
        gpu_ctx.render(1, 4, [[]], cabi.RENDER_WHIRL_BYPASS, ids=[0], inputs=np.zeros((1, 512), np.float32))


def test_leslie_only_on_caller_input(ref, gpu_ctx):
    """TBF_RENDER_WHIRL_INPUT: whirlProc / whirlProc3 on the caller's mono buffers, the second
    half of the hosts' chain (src/clap.cpp:254-258).  Noise with awkward values, rotor events,
    two batches back to back, against the reference's whirlProc3 / whirlProc."""
    rng = np.random.RandomState(7)
    n_frag = 220
    for flags in (cabi.RENDER_WHIRLPROC3, 0):
        r, first = _make(ref, gpu_ctx, 3, seed=2)
        x = (rng.standard_normal((3, n_frag * 128)) * 0.3).astype(np.float32)
        x[0, 1000:1400] = 0.0
        x[1, 5000] = np.float32(1e-38)          # denormal region, DENORMAL_HACK path
        x[1, 5001] = np.float32(-3.5)
        x[2, :128] = 0.0
        ev = [(0, cabi.EV_REV_OPTION, 8, 0), (60, cabi.EV_REV_SELECT, 1, 0), (130, cabi.EV_REV_OPTION, 4, 0),
              (131, cabi.EV_KEY_ON, 60, 0)]     # key events are not the Leslie's business here
        cut = 97
        ev_a = [e for e in ev if e[0] < cut]
        ev_b = [(f - cut, t, a, b) for (f, t, a, b) in ev if f >= cut]
        ids = list(range(first, first + 3))
        la, ra = gpu_ctx.render(3, cut, [ev_a] * 3, flags, ids=ids, inputs=x[:, :cut * 128])
        lb, rb = gpu_ctx.render(3, n_frag - cut, [ev_b] * 3, flags, ids=ids, inputs=x[:, cut * 128:])
        gl, gr = np.concatenate([la, lb], axis=1), np.concatenate([ra, rb], axis=1)
        for i in range(3):
            ri = ref.RefInstance(seed=2)
            rl, rr = ri.whirl_only(n_frag, flags, x[i], ev)
            ri.close()
            assert np.array_equal(_bits(gl[i]), _bits(rl)) and np.array_equal(_bits(gr[i]), _bits(rr)), (flags, i)
        # the tone generator was left alone: no key is down, nothing is listed
        assert gpu_ctx.export_state(first).active_osc_l_end == 0
        r.close()


def test_two_halves_with_host_audio_in_between(ref, gpu_ctx):
    """tonegen on the GPU (WHIRL_BYPASS) -> audio on the host (where a host would run its
    preamp / reverb) -> Leslie on the GPU (WHIRL_INPUT) == the contiguous chain, bit for bit."""
    n_frag = 260
    ev = workloads.instance_events(5, n_frag, config=3)
    r, first = _make(ref, gpu_ctx, 1)
    rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, ev)
    r.close()
    rotor = (cabi.EV_REV_OPTION, cabi.EV_REV_SELECT)
    mono, _ = gpu_ctx.render(1, n_frag, [[e for e in ev if e[1] not in rotor]],
                             cabi.RENDER_WHIRL_BYPASS, ids=[first])
    gl, gr = gpu_ctx.render(1, n_frag, [[e for e in ev if e[1] in rotor]], cabi.RENDER_WHIRLPROC3,
                            ids=[first], inputs=mono)
    assert np.array_equal(_bits(gl[0]), _bits(rl)) and np.array_equal(_bits(gr[0]), _bits(rr))
