"""MIDI front end (include/tbf_midi.h): raw MIDI -> event lists.

Pins, in the order the reference's own tests give them (src/midi.cpp:1460-1873):
  * the key-table known answers of its doctest cases (initMidiTables, transposes, splits),
  * the mapping observed on the reference itself (oracle: parse_raw_midi_data on a live engine),
  * equivalence of the whole translation: the oracle fed RAW MIDI renders the same bits as the
    oracle fed our compiled events (CPU), and as the CUDA path fed them (gpu).
"""
import math
import numpy as np
import pytest

from tunebfree_b200 import cabi, midi


def _bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


# ---------------------------------------------------------------- key tables


def test_function_table_follows_the_reference_names_and_prefix_rule():
    names = midi.function_names()
    assert len(names) == 69 and names[0] == "upper.drawbar16" and names[-1] == "convolution.mix"
    assert names[27:31] == ["percussion.enable", "percussion.decay", "percussion.harmonic",
                            "percussion.volume"]
    # getCCFunctionId: first table entry that is a prefix of the name (src/midi.cpp:333-346)
    assert midi.function_id("upper.drawbar1") == 8
    assert midi.function_id("upper.drawbar16") == 0
    assert midi.function_id("upper.drawbar135") == 6
    assert midi.function_id("upper.drawbar113-") == 7        # trailing '-' = inverted
    assert names[midi.function_id("xov.ctl_biased_fb2")] == "xov.ctl_biased_fb2"
    assert names[midi.function_id("xov.ctl_biased_fb")] == "xov.ctl_biased_fb"
    assert names[midi.function_id("xov.ctl_biased")] == "xov.ctl_biased"
    assert midi.function_id("no.such.function") == -1


def test_default_key_tables_reference_kats():
    """'Testing initMidiTables' and 'Testing map_to_real_key' (src/midi.cpp:1485-1518, :1594)."""
    m = midi.MidiMap()
    for n in (0, 35, 36, 37, 95, 96, 97, 127):
        assert m.key(0, n) == n
        assert m.key(1, n) == 128 + n
    for n in (0, 23, 24, 25, 54, 55, 56, 127):
        assert m.key(2, n) == 256 + n
    assert all(m.key(ch, 60) == -1 for ch in range(3, 16))


@pytest.mark.parametrize("t", [1, -1])
def test_transpose_reference_kats(t):
    """'Testing setKeyboardTranspose' (+ negative), src/midi.cpp:1520-1592."""
    m = midi.MidiMap(cfg="midi.transpose=%d\n" % t)
    for n in (0, 35, 36, 37, 95, 96, 97, 127):
        assert m.key(0, n) == n + t
        assert m.key(1, n) == 128 + n + t
    for n in (0, 23, 24, 25, 54, 55, 56, 127):
        assert m.key(2, n) == 256 + n + t


def test_channel_transposes_reference_kats():
    """'Testing loadKeyTable{A,B,C} with channel transpose' (src/midi.cpp:1660-1700)."""
    m = midi.MidiMap(cfg="midi.upper.transpose=3\nmidi.lower.transpose=-2\nmidi.pedals.transpose=5\n")
    assert [m.key(0, n) for n in (0, 60, 127)] == [3, 63, 130]
    assert [m.key(1, n) for n in (0, 60, 127)] == [126, 186, 253]
    assert [m.key(2, n) for n in (0, 24, 100)] == [261, 285, 361]


def test_splits_reference_kats():
    """'Testing loadKeyTableA with splits' / '... split transposes' (src/midi.cpp:1629-1658,
    :1702-1736): pedals below 40, lower manual below 66, upper manual above."""
    m = midi.MidiMap(split=(3, 40, 66, 0, 0, 0))
    assert [m.key(0, n) for n in (0, 23, 24, 25, 38, 39)] == [256 + n for n in (0, 23, 24, 25, 38, 39)]
    assert [m.key(0, n) for n in (40, 41, 65)] == [128 + 40, 128 + 41, 128 + 65]
    assert [m.key(0, n) for n in (66, 95, 96, 97, 127)] == [66, 95, 96, 97, 127]
    m = midi.MidiMap(split=(31, 40, 66, 3, 4, 2))
    assert [m.key(0, n) for n in (0, 39)] == [256 + 3, 256 + 39 + 3]
    assert [m.key(0, n) for n in (40, 65)] == [128 + 40 + 4, 128 + 65 + 4]
    assert [m.key(0, n) for n in (66, 127)] == [66 + 2, 127 + 2]


CFGS = [
    "",
    "midi.transpose=-12\nmidi.upper.transpose=2\nmidi.lower.transpose=-3\nmidi.pedals.transpose=7\n",
    "midi.upper.channel=3\nmidi.lower.channel=1\nmidi.pedals.channel=2\n",
    "midi.upper.channel=2\nmidi.lower.channel=2\n",     # shared channel: the lower manual wins
]


@pytest.mark.parametrize("cfg", CFGS)
def test_key_mapping_matches_the_reference_engine(ref, cfg):
    """Every (channel, note) probed on the live reference: parse_raw_midi_data() -> activeKeys[]."""
    r = ref.RefInstance(seed=1, cfg=cfg)
    m = midi.MidiMap(cfg=cfg)
    for ch in range(16):
        for note in (0, 1, 11, 12, 13, 36, 59, 60, 61, 96, 115, 126, 127):
            want = r.midi_key(ch, note)
            got = m.key(ch, note)
            got = got if 0 <= got < midi.MAX_KEYS else -1
            assert got == want, (cfg, ch, note)
    r.close()


# ---------------------------------------------------------------- controllers


def test_default_controller_map():
    """midiPrimeControllerMapping (src/midi.cpp:839-900): one controller per function and
    channel set, the last assignment wins (percussion.enable: 80 on the upper channel)."""
    m = midi.MidiMap()
    feet = ["16", "513", "8", "4", "223", "2", "135", "113", "1"]
    for ch, man in enumerate(("upper", "lower", "pedal")):
        assert [m.function(ch, 70 + i) for i in range(9)] == ["%s.drawbar%s" % (man, f) for f in feet]
        assert m.function(ch, 7) == "swellpedal1" and m.function(ch, 11) == "swellpedal2"
        assert m.function(ch, 1) == "rotary.speed-preset" and m.function(ch, 64) == "rotary.speed-toggle"
    assert m.function(0, 80) == "percussion.enable" and m.function(0, 66) is None
    assert m.function(1, 66) == "percussion.enable" and m.function(2, 66) == "percussion.enable"
    assert [m.function(0, c) for c in (81, 82, 83, 95, 92, 30, 31)] == [
        "percussion.volume", "percussion.decay", "percussion.harmonic", "vibrato.routing",
        "vibrato.knob", "vibrato.lower", "vibrato.upper"]
    assert m.function(1, 92) is None and m.function(3, 70) is None


def test_controller_value_quantisation():
    """setMIDIDrawBar: rint((127 - v) * 8 / 127) (src/tonegen.cpp:2752-2756); switches at 64;
    vibrato knob v / 23, routing v / 32; rotary preset v / 43 (src/whirl.cpp:244-249)."""
    m = midi.MidiMap()
    for v in range(128):
        ev, _ = m.compile([(0, bytes([0xB0, 72, v]))])
        assert ev == [(0, cabi.EV_DRAWBAR, 2, int(np.rint((127 - v) * 8.0 / 127.0)))]
        ev, _ = m.compile([(0, bytes([0xB1, 72, v]))])
        assert ev[0][2] == 11
        ev, _ = m.compile([(0, bytes([0xB0, 80, v])), (0, bytes([0xB0, 82, v])),
                           (0, bytes([0xB0, 83, v])), (0, bytes([0xB0, 81, v]))])
        on = 1 if v >= 64 else 0
        assert [e[1:3] for e in ev] == [(cabi.EV_PERC_ENABLE, on), (cabi.EV_PERC_FAST, on),
                                        (cabi.EV_PERC_FIRST, on), (cabi.EV_PERC_SOFT, on)]
        ev, _ = m.compile([(0, bytes([0xB0, 92, v])), (0, bytes([0xB0, 95, v])), (0, bytes([0xB0, 1, v])),
                           (0, bytes([0xB0, 7, v]))])
        assert ev == [(0, cabi.EV_VIB_KNOB, v // 23, 0), (0, cabi.EV_VIB_UPPER, (v // 32) >> 1, 0),
                      (0, cabi.EV_VIB_LOWER, (v // 32) & 1, 0), (0, cabi.EV_REV_SELECT, v // 43, 0),
                      (0, cabi.EV_SWELL, v, 0)]


def test_messages_without_an_event_are_counted():
    m = midi.MidiMap()
    ev, st = m.compile([(0, b"\xe0\x00\x40"), (0, b"\xb0\x00\x01"), (0, b"\xb0\x20\x01"),
                        (0, b"\xb0\x79\x00"), (0, b"\xb0\x7e\x00"), (0, b"\xb0\x5b\x40"),
                        (0, b"\xb0\x03\x40"), (0, b"\x95\x3c\x64"), (0, b"\x90\x3c"),
                        (0, b"\xc0\x05"), (1, b"\xb0\x40\x20")])
    assert ev == [] and st.n_ignored == 9 and st.n_unsupported == 1  # pedal release: no action
    m = midi.MidiMap(cfg="midi.controller.upper.20=whirl.horn.filter.a.hz\nmidi.transpose=100\n")
    ev, st = m.compile([(0, b"\xb0\x14\x40"), (0, b"\x92\x7f\x64"), (0, b"\x90\x00\x64")])
    assert [e[1] for e in ev] == [cabi.EV_HORN_FILTER] * 5 + [cabi.EV_KEY_ON]
    assert ev[-1] == (0, cabi.EV_KEY_ON, 100, 0) and st.n_unsupported == 0 and st.n_unmapped == 1
    ev, _ = midi.MidiMap().compile([(3, b"\xb0\x7b\x00")])
    assert [e[2] for e in ev] == list(range(384)) and all(e[1] == cabi.EV_KEY_OFF for e in ev)


# ---------------------------------------------------------------- whole translation


def midi_script(seed, n_fragments, channels=(0, 1, 2), extra_cc=(), dense_cc=()):
    """A seeded performance: chords on the three manuals, drawbar / percussion / vibrato /
    rotary / swell controllers, the sustain pedal, one panic.  dense_cc: (status, controller)
    pairs that get a random value every few fragments."""
    rng = np.random.RandomState(seed)
    rng2 = np.random.RandomState(seed + 7919)
    up, lo, pd = channels
    msgs = [(0, bytes([0xB0 | up, 70 + i, 0])) for i in range(9)]            # 888888888
    msgs += [(0, bytes([0xB0 | lo, 70 + i, 40])) for i in range(4)]
    msgs += [(0, bytes([0xB0 | pd, 70, 10])), (0, bytes([0xB0 | pd, 72, 30]))]
    msgs += [(0, bytes([0xB0 | up, 80, 127])), (0, bytes([0xB0 | up, 95, 127])),
             (0, bytes([0xB0 | up, 92, 120]))]
    held = []
    f = 1
    while f < n_fragments:
        kind = rng.randint(0, 12)
        if kind < 5 or not held:
            ch = (up, up, lo, pd)[rng.randint(0, 4)]
            note = int(rng.randint(24, 50) if ch == pd else rng.randint(36, 97))
            msgs.append((f, bytes([0x90 | ch, note, int(rng.randint(1, 128))])))
            held.append((ch, note))
        elif kind < 8:
            ch, note = held.pop(rng.randint(0, len(held)))
            off = bytes([0x80 | ch, note, 0]) if rng.randint(0, 2) else bytes([0x90 | ch, note, 0])
            msgs.append((f, off))
        elif kind == 8:
            ch = (up, lo, pd)[rng.randint(0, 3)]
            msgs.append((f, bytes([0xB0 | ch, 70 + int(rng.randint(0, 9)), int(rng.randint(0, 128))])))
        elif kind == 9:
            cc = (64, 64, 1, 7, 11)[rng.randint(0, 5)]
            msgs.append((f, bytes([0xB0 | up, cc, int(rng.randint(0, 128))])))
        elif kind == 10:
            cc = (80, 81, 82, 83, 92, 95, 30, 31)[rng.randint(0, 8)]
            msgs.append((f, bytes([0xB0 | up, cc, int(rng.randint(0, 128))])))
        else:
            pool = [(0xB0 | up, 91, 64), (0xE0 | up, 0, 64), (0xB0 | lo, 66, 127), (0xB0 | 9, 70, 0)]
            pool += list(extra_cc)
            msgs.append((f, bytes(pool[rng.randint(0, len(pool))])))
        if dense_cc and rng2.randint(0, 2):
            st, cc = dense_cc[rng2.randint(0, len(dense_cc))]
            # the ends of the range are where setIIRFilter() rejects Q and gain
            v = int(rng2.randint(0, 128)) if rng2.randint(0, 4) else (0, 1, 2, 126, 127)[rng2.randint(0, 5)]
            msgs.append((f, bytes([st, cc, v])))
        f += int(rng.randint(1, 12))
    msgs.append((n_fragments - 20, bytes([0xB0 | up, 123, 0])))
    return sorted(msgs, key=lambda t: t[0])


WHIRL_CC = ["whirl.horn.filter.a.type", "whirl.horn.filter.a.hz", "whirl.horn.filter.a.q",
            "whirl.horn.filter.a.gain", "whirl.horn.filter.b.type", "whirl.horn.filter.b.hz",
            "whirl.horn.filter.b.q", "whirl.horn.filter.b.gain", "whirl.horn.brakepos",
            "whirl.drum.brakepos", "whirl.horn.acceleration", "whirl.horn.deceleration",
            "whirl.drum.acceleration", "whirl.drum.deceleration"]

WHIRL_CC_NUM = list(range(20, 32)) + [34, 35]   # 32 is bank select LSB: never dispatched

SCRIPT_CFGS = [
    ("", (0, 1, 2), ()),
    # (receive channels stay within 1..3: above, the reference's own oscKeyOn writes its
    #  _activeKeys bitmap out of bounds, src/tonegen.cpp:3150 with src/midi.cpp:1079-1083)
    ("midi.transpose=-5\nmidi.lower.transpose=4\nmidi.upper.channel=3\nmidi.pedals.channel=1\n"
     "midi.controller.upper.20=upper.drawbar8-\nmidi.controller.upper.64=unmap\n"
     "midi.controller.upper.21=rotary.speed-select\nmidi.controller.lower.80=percussion.enable\n",
     (2, 1, 0), ((0xB2, 20, 90), (0xB2, 21, 100), (0xB2, 21, 33), (0xB1, 80, 10))),
    ("whirl.speed-preset=2\nmidi.controller.reset=1\nmidi.controller.upper.64=rotary.speed-toggle\n"
     "midi.controller.upper.70=vibrato.knob\nmidi.controller.pedals.7=swellpedal1\n",
     (0, 1, 2), ((0xB0, 64, 127), (0xB2, 7, 99), (0xB0, 70, 77))),
    # the Leslie's run-time controllers (src/whirl.cpp:966-981): horn filters, brakes, ramps
    ("".join("midi.controller.upper.%d=%s\n" % (c, n) for c, n in zip(WHIRL_CC_NUM, WHIRL_CC)),
     (0, 1, 2), ((0xB0, 1, 0), (0xB0, 1, 50), (0xB0, 1, 100)), tuple((0xB0, c) for c in WHIRL_CC_NUM)),
]


@pytest.mark.parametrize("case", range(len(SCRIPT_CFGS)))
def test_compiled_events_render_like_raw_midi_on_the_reference(ref, case):
    """CPU: the reference engine fed raw MIDI == the reference engine fed our event list."""
    cfg, channels, extra = SCRIPT_CFGS[case][:3]
    msgs = midi_script(10 + case, 260, channels, extra, *SCRIPT_CFGS[case][3:])
    a = ref.RefInstance(seed=3, cfg=cfg)
    wt = a.whirl_tables()
    m = midi.MidiMap(cfg=cfg, revselects=list(wt.revselects))
    events, st = m.compile(msgs, whirl_tables=wt)
    assert st.n_unsupported == 0 and len(events) > 100
    al, ar = a.render_midi(260, cabi.RENDER_WHIRLPROC3, msgs)
    a.close()
    b = ref.RefInstance(seed=3, cfg=cfg)
    bl, br = b.render(260, cabi.RENDER_WHIRLPROC3, events)
    b.close()
    assert np.abs(al).max() > 0.01
    assert np.array_equal(_bits(al), _bits(bl)) and np.array_equal(_bits(ar), _bits(br))


def test_leslie_controllers_leave_the_reference_in_the_compiled_state(ref):
    """Every value of every whirl.* controller, one after the other, through the reference's
    parse_raw_midi_data(); the coefficients / brake positions / ramp factors its struct b_whirl
    ends up with == what the compiled events carry (filters: tbf_iir_coefficients on the host,
    incl. the values setIIRFilter() rejects)."""
    cfg = SCRIPT_CFGS[3][0]
    r = ref.RefInstance(seed=1, cfg=cfg)
    wt = r.whirl_tables()
    m = midi.MidiMap(cfg=cfg, revselects=list(wt.revselects))
    st = m.new_state(wt)
    rng = np.random.RandomState(5)
    order = [(i, v) for i in range(len(WHIRL_CC)) for v in range(128)]
    rng.shuffle(order)
    iir = [list(wt.hafw), list(wt.hbfw)]
    brake = [wt.hn_brake_pos, wt.dr_brake_pos]
    ramp_v = [None] * 4
    f32 = lambda x: np.asarray(x, np.float32).view(np.uint32)
    for start in range(0, len(order), 64):
        msgs = [(0, bytes([0xB0, WHIRL_CC_NUM[i], v])) for i, v in order[start:start + 64]]
        ev, st = m.compile(msgs, state=st)
        for _, t, a, b in ev:
            if t == cabi.EV_HORN_FILTER:
                iir[a >> 3][a & 7] = np.int32(b).view(np.float32)
            elif t == cabi.EV_WHIRL_BRAKE:
                brake[a] = b / 127.0
            else:
                assert t == cabi.EV_WHIRL_RAMP
                ramp_v[a] = b
        r.render_midi(1, 0, msgs)
        rs = r.export_state()
        assert np.array_equal(f32(iir), f32(np.ctypeslib.as_array(rs.horn_iir))), start
        assert tuple(rs.brake_pos) == tuple(brake)
        for k in range(4):
            if ramp_v[k] is not None:
                tau = np.float32(.01 + ramp_v[k] / (80.0 if k < 2 else 14.0))
                assert rs.ramp_one_minus_l[k] == 1 - math.exp(-1.0 / (48000.0 / 128.0 * float(tau)))
    assert st.n_unsupported == 0 and st.n_ignored == 0
    r.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", [1, 3])
def test_gpu_render_of_compiled_midi_matches_the_reference_fed_raw_midi(ref, gpu_ctx, case):
    """GPU: 6 organs, each with its own MIDI performance, against parse_raw_midi_data()."""
    cfg, channels, extra = SCRIPT_CFGS[case][:3]
    n_frag = 300
    r0 = ref.RefInstance(seed=4, cfg=cfg)
    wt = r0.whirl_tables()
    tg = gpu_ctx.tonegen_tables(r0.tonegen_tables())
    wh = gpu_ctx.whirl_tables(wt)
    r0.close()
    n = 6
    first = gpu_ctx.create_instances(tg, wh, n)
    m = midi.MidiMap(cfg=cfg, revselects=list(wt.revselects))
    scripts = [midi_script(100 + i, n_frag, channels, extra, *SCRIPT_CFGS[case][3:]) for i in range(n)]
    events = [m.compile(s, whirl_tables=wt)[0] for s in scripts]
    gl, gr = gpu_ctx.render(n, n_frag, events, cabi.RENDER_WHIRLPROC3,
                            ids=list(range(first, first + n)))
    for i in range(n):
        r = ref.RefInstance(seed=4, cfg=cfg)
        rl, rr = r.render_midi(n_frag, cabi.RENDER_WHIRLPROC3, scripts[i])
        r.close()
        assert np.array_equal(_bits(gl[i]), _bits(rl)) and np.array_equal(_bits(gr[i]), _bits(rr)), i


# ---------------------------------------------------------------- programmes

PGM_TEXT = """
# programmes of the test (syntax: src/pgmParser.cpp:65-75)
1 { name="Full organ", drawbars="88 8888 888", drawbarslower = 00-8400-000
    drawbarspedals=80_0000_000 vibrato=c3 vibratoupper=on vibratolower=off
    perc=on percvol=soft percspeed=fast percharm=second rotaryspeed=tremolo reverbmix=0.2 }
2 { name=Jazz drawbars=888000000, perc=yes, percvol=hi, percspeed=slow, percharm=3rd,
    vibratoknob=v1, vibratoupper=no, overdrive=off, rotaryspeed=chorale }
3 {name="Split \\"A\\"" keysplitlower=60 keysplitpedals=40 trssplitpedals=-12 trssplitlower=12
   trssplitupper=-5 vibratolower=yes rotaryspeed=stop}
7 { transpose=2 transposeupper=-3 transposelower=4 transposepedals=1 perc=off drawbars=008800000 }
9 { name=random drawbars=random }
"""


def test_programme_text_parser():
    m = midi.MidiMap(programmes=PGM_TEXT)
    t = m.programmes
    g = t.prog[1]
    assert g.name == b"Full organ" and list(g.drawbars[0]) == [8] * 9
    assert list(g.drawbars[1]) == [0, 0, 8, 4, 0, 0, 0, 0, 0] and list(g.drawbars[2]) == [8] + [0] * 8
    assert g.scanner == (0x83 | 0x200) and (g.perc_enabled, g.perc_soft, g.perc_fast, g.perc_second) == (1, 1, 1, 1)
    assert g.rotary_speed == 2 and abs(g.reverb_mix - 0.2) < 1e-7
    inuse, drawbr, lowdrw, pdldrw, scannr = 0x1, 0x2, 0x80000, 0x100000, 0x100
    assert g.flags & (inuse | drawbr | lowdrw | pdldrw | scannr) == (inuse | drawbr | lowdrw | pdldrw | scannr)
    assert t.prog[2].name == b"Jazz" and t.prog[2].scanner & 0xFF == 1 and t.prog[2].rotary_speed == 0
    assert t.prog[3].name == b'Split "A"' and (t.prog[3].split_lower, t.prog[3].split_pedals) == (60, 40)
    assert list(t.prog[7].transpose)[:4] == [2, -3, 4, 1] and t.prog[7].perc_enabled == 0
    assert t.prog[9].flags & 0x20000 and t.prog[4].flags == 0
    for bad, line in (("1 { drawbars=123 }", 1), ("1 {\n perc=maybe }", 2), ("\n\n5 { nosuch=1 }", 3),
                      ("1 { vibrato=v9 }", 1), ("x { }", 1), ("1 { transpose=200 }", 1), ("1 { name= }", 1)):
        with pytest.raises(ValueError, match="line %d" % line):
            midi.MidiMap(programmes=bad)


def test_programme_change_events():
    m = midi.MidiMap(programmes=PGM_TEXT)
    ev, st = m.compile([(5, b"\xc0\x00")])          # programme 0 + offset 1
    want = [(cabi.EV_DRAWBAR, i, 8) for i in range(9)]
    want += [(cabi.EV_DRAWBAR, 9 + i, v) for i, v in enumerate([0, 0, 8, 4, 0, 0, 0, 0, 0])]
    want += [(cabi.EV_DRAWBAR, 18 + i, v) for i, v in enumerate([8] + [0] * 8)]
    want += [(cabi.EV_VIB_KNOB, 5, 0), (cabi.EV_VIB_UPPER, 1, 0), (cabi.EV_VIB_LOWER, 0, 0),
             (cabi.EV_VIB_UPPER, 1, 0), (cabi.EV_VIB_LOWER, 0, 0),
             (cabi.EV_PERC_ENABLE, 1, 0), (cabi.EV_PERC_ENABLE, 1, 0), (cabi.EV_PERC_SOFT, 1, 0),
             (cabi.EV_PERC_FAST, 1, 0), (cabi.EV_PERC_FIRST, 1, 0), (cabi.EV_REV_SELECT, 2, 0)]
    assert [e[1:] for e in ev] == want and all(e[0] == 5 for e in ev)
    assert (st.vib_routing, st.rev_select, st.n_unsupported) == (2, 2, 0)
    # splits and transposes move the instance's keys, not the shared map
    st = m.new_state()
    ev, st = m.compile([(0, b"\xc0\x02"), (1, b"\x90\x1e\x40"), (1, b"\x90\x32\x40"), (1, b"\x90\x46\x40")],
                       state=st)
    assert [e[1:3] for e in ev[-3:]] == [(cabi.EV_KEY_ON, 256 + 30 - 12), (cabi.EV_KEY_ON, 128 + 50 + 12),
                                         (cabi.EV_KEY_ON, 70 - 5)]
    assert m.key(0, 30) == 30 and m.key_now(st, 0, 30) == 256 + 18
    ev, st = m.compile([(0, b"\xc0\x06"), (1, b"\x90\x3c\x40"), (1, b"\x91\x3c\x40"), (1, b"\x92\x18\x40")])
    assert [e[2] for e in ev if e[1] == cabi.EV_KEY_ON] == [60 + 2 - 3, 128 + 60 + 2 + 4, 256 + 24 + 2 + 1]
    # unused numbers, the table end and random drawbars
    ev, st = m.compile([(0, b"\xc0\x03"), (0, b"\xc0\x7f"), (0, b"\xc0\x08")])
    assert ev == [] and (st.n_ignored, st.n_unsupported) == (2, 1)
    m0 = midi.MidiMap(programmes=PGM_TEXT, cfg="pgm.controller.offset=0\n")
    assert [e[3] for e in m0.compile([(0, b"\xc0\x02")])[0][:3]] == [8, 8, 8]


def _programme_script(seed, n_fragments):
    rng = np.random.RandomState(seed)
    msgs = midi_script(seed, n_fragments)
    for f in sorted(rng.randint(2, n_fragments - 30, size=14)):
        msgs.append((int(f), bytes([0xC0 | int(rng.randint(0, 3)), int(rng.choice([0, 1, 2, 6, 6, 1, 2]))])))
    return sorted(msgs, key=lambda t: t[0])


def test_programme_changes_render_like_the_reference(ref):
    """CPU: the reference with the same programme text (its own parser) and raw MIDI, programme
    changes included, against the reference fed our events."""
    msgs = _programme_script(21, 300)
    a = ref.RefInstance(seed=2)
    a.load_programmes(PGM_TEXT)
    wt = a.whirl_tables()
    m = midi.MidiMap(revselects=list(wt.revselects), programmes=PGM_TEXT)
    events, st = m.compile(msgs, whirl_tables=wt)
    assert st.n_unsupported == 0 and sum(1 for e in events if e[1] == cabi.EV_REV_SELECT) >= 3
    al, ar = a.render_midi(300, cabi.RENDER_WHIRLPROC3, msgs)
    a.close()
    b = ref.RefInstance(seed=2)
    bl, br = b.render(300, cabi.RENDER_WHIRLPROC3, events)
    b.close()
    assert np.abs(al).max() > 0.01
    assert np.array_equal(_bits(al), _bits(bl)) and np.array_equal(_bits(ar), _bits(br))


def test_reference_default_programme_file(ref):
    """pgm/default.pgm of the reference tree, when it is there (not on the GPU box): parsed by
    both sides, every programme installed once, same audio."""
    import os
    path = "/root/reference/pgm/default.pgm"
    if not os.path.exists(path):
        pytest.skip("reference tree not present")
    text = open(path).read()
    m = midi.MidiMap(programmes=text)
    used = [p for p in range(1, 129) if m.programmes.prog[p].flags & 1]
    assert len(used) > 100
    # programmes with drawbars=random draw from the process-global rand(): not reproducible
    used = [p for p in used if not m.programmes.prog[p].flags & 0x20000]
    msgs = [(0, b"\x90\x3c\x64"), (0, b"\x91\x30\x64"), (0, b"\x92\x1c\x64")]
    msgs += [(2 + 3 * i, bytes([0xC0, p - 1])) for i, p in enumerate(used)]
    n_frag = 2 + 3 * len(used) + 4
    a = ref.RefInstance(seed=1)
    a.load_programmes(text)
    wt = a.whirl_tables()
    m = midi.MidiMap(revselects=list(wt.revselects), programmes=text)
    events, st = m.compile(msgs, whirl_tables=wt)
    al, ar = a.render_midi(n_frag, cabi.RENDER_WHIRLPROC3, msgs)
    a.close()
    b = ref.RefInstance(seed=1)
    bl, br = b.render(n_frag, cabi.RENDER_WHIRLPROC3, events)
    b.close()
    assert st.n_unsupported == 0
    assert np.array_equal(_bits(al), _bits(bl)) and np.array_equal(_bits(ar), _bits(br))


@pytest.mark.gpu
def test_gpu_render_of_programme_changes_matches_the_reference(ref, gpu_ctx):
    """GPU: 4 organs playing through programme changes (drawbars, percussion, vibrato, rotary,
    splits and transposes) against the reference with the same programme text and raw MIDI."""
    n_frag, n = 300, 4
    r0 = ref.RefInstance(seed=6)
    wt = r0.whirl_tables()
    tg = gpu_ctx.tonegen_tables(r0.tonegen_tables())
    wh = gpu_ctx.whirl_tables(wt)
    r0.close()
    first = gpu_ctx.create_instances(tg, wh, n)
    m = midi.MidiMap(revselects=list(wt.revselects), programmes=PGM_TEXT)
    scripts = [_programme_script(40 + i, n_frag) for i in range(n)]
    events = [m.compile(s, whirl_tables=wt)[0] for s in scripts]
    gl, gr = gpu_ctx.render(n, n_frag, events, cabi.RENDER_WHIRLPROC3,
                            ids=list(range(first, first + n)))
    for i in range(n):
        r = ref.RefInstance(seed=6)
        r.load_programmes(PGM_TEXT)
        rl, rr = r.render_midi(n_frag, cabi.RENDER_WHIRLPROC3, scripts[i])
        r.close()
        assert np.array_equal(_bits(gl[i]), _bits(rl)) and np.array_equal(_bits(gr[i]), _bits(rr)), i
