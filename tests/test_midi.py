"""MIDI front end (include/tbf_midi.h): raw MIDI -> event lists.

Pins, in the order the reference's own tests give them (src/midi.cpp:1460-1873):
  * the key-table known answers of its doctest cases (initMidiTables, transposes, splits),
  * the mapping observed on the reference itself (oracle: parse_raw_midi_data on a live engine),
  * equivalence of the whole translation: the oracle fed RAW MIDI renders the same bits as the
    oracle fed our compiled events (CPU), and as the CUDA path fed them (gpu).
"""
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
    assert ev == [(0, cabi.EV_KEY_ON, 100, 0)] and st.n_unsupported == 1 and st.n_unmapped == 1
    ev, _ = midi.MidiMap().compile([(3, b"\xb0\x7b\x00")])
    assert [e[2] for e in ev] == list(range(384)) and all(e[1] == cabi.EV_KEY_OFF for e in ev)


# ---------------------------------------------------------------- whole translation


def midi_script(seed, n_fragments, channels=(0, 1, 2), extra_cc=()):
    """A seeded performance: chords on the three manuals, drawbar / percussion / vibrato /
    rotary / swell controllers, the sustain pedal, one panic."""
    rng = np.random.RandomState(seed)
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
        f += int(rng.randint(1, 12))
    msgs.append((n_fragments - 20, bytes([0xB0 | up, 123, 0])))
    return sorted(msgs, key=lambda t: t[0])


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
]


@pytest.mark.parametrize("case", range(len(SCRIPT_CFGS)))
def test_compiled_events_render_like_raw_midi_on_the_reference(ref, case):
    """CPU: the reference engine fed raw MIDI == the reference engine fed our event list."""
    cfg, channels, extra = SCRIPT_CFGS[case]
    msgs = midi_script(10 + case, 260, channels, extra)
    a = ref.RefInstance(seed=3, cfg=cfg)
    wt = a.whirl_tables()
    m = midi.MidiMap(cfg=cfg, revselects=list(wt.revselects))
    events, st = m.compile(msgs, init_rev_select=wt.init_rev_select)
    assert st.n_unsupported == 0 and len(events) > 100
    al, ar = a.render_midi(260, cabi.RENDER_WHIRLPROC3, msgs)
    a.close()
    b = ref.RefInstance(seed=3, cfg=cfg)
    bl, br = b.render(260, cabi.RENDER_WHIRLPROC3, events)
    b.close()
    assert np.abs(al).max() > 0.01
    assert np.array_equal(_bits(al), _bits(bl)) and np.array_equal(_bits(ar), _bits(br))


@pytest.mark.gpu
def test_gpu_render_of_compiled_midi_matches_the_reference_fed_raw_midi(ref, gpu_ctx):
    """GPU: 6 organs, each with its own MIDI performance, against parse_raw_midi_data()."""
    cfg, channels, extra = SCRIPT_CFGS[1]
    n_frag = 300
    r0 = ref.RefInstance(seed=4, cfg=cfg)
    wt = r0.whirl_tables()
    tg = gpu_ctx.tonegen_tables(r0.tonegen_tables())
    wh = gpu_ctx.whirl_tables(wt)
    r0.close()
    n = 6
    first = gpu_ctx.create_instances(tg, wh, n)
    m = midi.MidiMap(cfg=cfg, revselects=list(wt.revselects))
    scripts = [midi_script(100 + i, n_frag, channels, extra) for i in range(n)]
    events = [m.compile(s, init_rev_select=wt.init_rev_select)[0] for s in scripts]
    gl, gr = gpu_ctx.render(n, n_frag, events, cabi.RENDER_WHIRLPROC3,
                            ids=list(range(first, first + n)))
    for i in range(n):
        r = ref.RefInstance(seed=4, cfg=cfg)
        rl, rr = r.render_midi(n_frag, cabi.RENDER_WHIRLPROC3, scripts[i])
        r.close()
        assert np.array_equal(_bits(gl[i]), _bits(rl)) and np.array_equal(_bits(gr[i]), _bits(rr)), i
