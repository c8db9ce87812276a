"""Raw MIDI -> event lists (include/tbf_midi.h): the host-side front end that stands where
parse_raw_midi_data() / process_midi_event() stand in the reference (src/midi.cpp:1095-1302).

    m = MidiMap(cfg="midi.transpose=-12\\n", revselects=(4, 0, 8))
    events, state = m.compile([(0, b"\\x90\\x3c\\x64"), (375, b"\\x80\\x3c\\x00")], instance=0)

`events` are (fragment, type, a, b) tuples in arrival order, ready for Context.render().
"""
import ctypes as C

from . import cabi
from .engine import lib as _gpu_lib
from .tables import parse_cfg_lines

MAX_FN = 80
MAX_KEYS = 384


class _Map(C.Structure):
    _fields_ = [
        ("rcv_ch", C.c_uint8 * 3),
        ("pad_", C.c_uint8),
        ("transpose", C.c_int32),
        ("nsh_a", C.c_int32), ("nsh_a_u", C.c_int32), ("nsh_a_pl", C.c_int32),
        ("nsh_a_ul", C.c_int32), ("nsh_b", C.c_int32), ("nsh_c", C.c_int32),
        ("split_a_pl", C.c_int32), ("split_a_ul", C.c_int32),
        ("cc_use", (C.c_uint8 * MAX_FN) * 3),
        ("cc_flags", (C.c_uint8 * 128) * 16),
        ("revselects", C.c_int32 * 3),
        ("key_table", (C.c_int16 * 128) * 16),
        ("cc_fn", (C.c_int16 * 128) * 16),
        ("ch_set", C.c_int8 * 16),
        ("programmes", C.c_void_p),
    ]


class Programme(C.Structure):
    _fields_ = [
        ("flags", C.c_uint32),
        ("drawbars", (C.c_uint8 * 9) * 3),
        ("perc_enabled", C.c_uint8), ("perc_soft", C.c_uint8), ("perc_fast", C.c_uint8),
        ("perc_second", C.c_uint8), ("overdrive", C.c_uint8),
        ("scanner", C.c_int16),
        ("rotary_speed", C.c_int16),
        ("split_lower", C.c_int16), ("split_pedals", C.c_int16),
        ("transpose", C.c_int16 * 7),
        ("reverb_mix", C.c_float),
        ("name", C.c_char * 22),
    ]


class ProgrammeTable(C.Structure):
    _fields_ = [
        ("controller_offset", C.c_int32),
        ("previous", C.c_int32),
        ("prog", Programme * 129),
    ]


class MidiState(C.Structure):
    _fields_ = [
        ("rev_select", C.c_int32),
        ("vib_routing", C.c_int32),
        ("transpose", C.c_int32), ("nsh_a", C.c_int32), ("nsh_a_u", C.c_int32),
        ("nsh_a_pl", C.c_int32), ("nsh_a_ul", C.c_int32), ("nsh_b", C.c_int32),
        ("nsh_c", C.c_int32), ("split_a_pl", C.c_int32), ("split_a_ul", C.c_int32),
        ("horn_filter", (C.c_float * 4) * 2),
        ("sample_rate", C.c_double),
        ("n_ignored", C.c_uint32),
        ("n_unsupported", C.c_uint32),
        ("n_unmapped", C.c_uint32),
    ]


_bound = False


def _lib():
    global _bound
    L = _gpu_lib()
    if not _bound:
        L.tbf_midi_map_default.argtypes = [C.POINTER(_Map)]
        L.tbf_midi_map_default.restype = None
        L.tbf_midi_map_apply_cfg.argtypes = [C.POINTER(_Map), C.c_char_p, C.c_char_p]
        L.tbf_midi_map_split.argtypes = [C.POINTER(_Map)] + [C.c_int] * 6
        L.tbf_midi_map_split.restype = None
        L.tbf_midi_map_finish.argtypes = [C.POINTER(_Map), C.POINTER(C.c_int32)]
        L.tbf_midi_state_init.argtypes = [C.POINTER(MidiState), C.POINTER(_Map),
                                          C.POINTER(cabi.WhirlTables)]
        L.tbf_midi_state_init.restype = None
        L.tbf_midi_key.argtypes = [C.POINTER(_Map), C.POINTER(MidiState), C.c_int, C.c_int]
        L.tbf_programmes_init.argtypes = [C.POINTER(ProgrammeTable)]
        L.tbf_programmes_init.restype = None
        L.tbf_programmes_load.argtypes = [C.POINTER(ProgrammeTable), C.c_char_p, C.POINTER(C.c_int)]
        L.tbf_programme_bind.argtypes = [C.POINTER(ProgrammeTable), C.c_int, C.c_char_p, C.c_char_p]
        L.tbf_midi_compile.argtypes = [C.POINTER(_Map), C.POINTER(MidiState), C.c_uint32, C.c_uint32,
                                       C.c_char_p, C.c_size_t, C.POINTER(cabi.Event), C.c_size_t]
        L.tbf_midi_function_name.argtypes = [C.c_int]
        L.tbf_midi_function_name.restype = C.c_char_p
        L.tbf_midi_function_id.argtypes = [C.c_char_p]
        _bound = True
    return L


def function_names():
    L = _lib()
    return [L.tbf_midi_function_name(i).decode() for i in range(L.tbf_midi_function_count())]


def function_id(name):
    return _lib().tbf_midi_function_id(name.encode())


class MidiMap:
    """Receive channels, key tables and controller assignments of one configuration."""

    def __init__(self, cfg="", revselects=None, split=None, programmes=None):
        """programmes: text in the syntax of a .pgm file (src/pgmParser.cpp), served on
        programme-change messages; None = programme changes are counted as unsupported."""
        L = _lib()
        self.c = _Map()
        L.tbf_midi_map_default(C.byref(self.c))
        self.programmes = None
        if programmes is not None:
            self.programmes = ProgrammeTable()
            L.tbf_programmes_init(C.byref(self.programmes))
            line = C.c_int(0)
            if L.tbf_programmes_load(C.byref(self.programmes), programmes.encode(), C.byref(line)) != 0:
                raise ValueError("programme text: error on line %d" % line.value)
            self.c.programmes = C.addressof(self.programmes)
        for key, value in parse_cfg_lines(cfg):
            if key.lower() == "pgm.controller.offset" and self.programmes is not None:
                if value.strip() in ("0", "1"):          # pgmConfig, src/program.cpp:926-937
                    self.programmes.controller_offset = int(value)
                continue
            rc = L.tbf_midi_map_apply_cfg(C.byref(self.c), key.encode(), value.encode())
            if rc < 0:
                raise ValueError("bad MIDI configuration line %s=%s" % (key, value))
        if split:  # (flags, split_a_pl, split_a_ul, nsh_a_pl, nsh_a_ul, nsh_a_u)
            L.tbf_midi_map_split(C.byref(self.c), *split)
        rs = (C.c_int32 * 3)(*revselects) if revselects is not None else None
        if L.tbf_midi_map_finish(C.byref(self.c), rs) != 0:
            raise ValueError("tbf_midi_map_finish failed")

    def key(self, channel, note):
        return int(self.c.key_table[channel][note])

    def function(self, channel, controller):
        f = int(self.c.cc_fn[channel][controller])
        return None if f < 0 else _lib().tbf_midi_function_name(f).decode()

    def new_state(self, whirl_tables=None):
        """State of a fresh instance of `whirl_tables` (cabi.WhirlTables; None = the defaults)."""
        st = MidiState()
        wt = C.byref(whirl_tables) if whirl_tables is not None else None
        _lib().tbf_midi_state_init(C.byref(st), C.byref(self.c), wt)
        return st

    def key_now(self, state, channel, note):
        """Engine key under the instance's CURRENT mapping (programmes may have moved it)."""
        return _lib().tbf_midi_key(C.byref(self.c), C.byref(state), channel, note)

    def compile(self, messages, instance=0, whirl_tables=None, state=None):
        """messages: iterable of (fragment, bytes) sorted by fragment -> (events, MidiState)."""
        L = _lib()
        st = state if state is not None else self.new_state(whirl_tables)
        buf = (cabi.Event * MAX_KEYS)()
        out = []
        for fragment, raw in messages:
            raw = bytes(raw)
            n = L.tbf_midi_compile(C.byref(self.c), C.byref(st), instance, fragment, raw, len(raw),
                                   buf, MAX_KEYS)
            if n < 0:
                raise ValueError("tbf_midi_compile failed (%d)" % n)
            out.extend((int(e.fragment), int(e.type), int(e.a), int(e.b)) for e in buf[:n])
        return out, st
