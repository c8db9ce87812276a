"""ORACLE (test infrastructure, not product code).

ctypes wrapper around ``oracle/_ref/libtbf_ref.so`` = the UNMODIFIED reference
engine compiled in place from /root/reference/src (see oracle/Makefile) plus
the harness ``oracle/ref_harness.cpp``.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from tunebfree_b200 import cabi

_HERE = os.path.dirname(os.path.abspath(__file__))
# TBF_REF_BUILD=stock: the engine compiled with the project's own -O3 -ffast-math flags, for
# bench.py's second timed CPU row only -- its output is NOT the parity oracle's
LIB_PATH = os.path.join(_HERE, "_ref", "libtbf_ref_stock.so" if os.environ.get("TBF_REF_BUILD") == "stock"
                        else "libtbf_ref.so")
_lib = None


def build(quiet=True):
    """Compile the oracle (needs /root/reference; a no-op when already built)."""
    if not os.path.isdir(os.environ.get("TBF_REFERENCE", "/root/reference")):
        if os.path.exists(LIB_PATH):
            return LIB_PATH
        raise RuntimeError("reference tree absent and oracle/_ref not prebuilt")
    subprocess.run(["make", "-C", _HERE, "-j8"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                 C.c_uint, C.c_char_p]
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_event.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.ref_render.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.POINTER(cabi.Event),
                                 C.c_ulonglong, C.c_void_p, C.c_void_p]
        L.ref_active_wheels.argtypes = [C.c_void_p]
        L.ref_render_midi.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.POINTER(C.c_uint),
                                      C.c_ulonglong, C.c_void_p, C.c_void_p]
        L.ref_midi_key.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_load_programmes.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_whirl_only.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p, C.POINTER(cabi.Event),
                                     C.c_ulonglong, C.c_void_p, C.c_void_p]
        L.ref_whirl2.argtypes = [C.c_void_p, C.c_uint] + [C.c_void_p] * 7
        L.ref_dump_tonegen_tables.argtypes = [C.c_void_p, C.POINTER(cabi.TonegenTables)]
        L.ref_free_tonegen_tables.argtypes = [C.POINTER(cabi.TonegenTables)]
        L.ref_dump_whirl_tables.argtypes = [C.c_void_p, C.POINTER(cabi.WhirlTables)]
        L.ref_export_state.argtypes = [C.c_void_p, C.POINTER(cabi.StateExport)]
        _lib = L
    return _lib


def tet12():
    """12-TET, A4 = 440 Hz: the table libMTSClient serves without a master."""
    return np.array([440.0 * 2.0 ** ((n - 69) / 12.0) for n in range(128)], dtype=np.float64)


def events_array(events):
    """[(fragment, type, a, b), ...] -> ctypes array of tbf_event (sorted by fragment)."""
    ev = sorted(events, key=lambda e: e[0])
    arr = (cabi.Event * max(1, len(ev)))()
    for i, e in enumerate(ev):
        arr[i].instance = 0
        arr[i].fragment = e[0]
        arr[i].type = e[1]
        arr[i].a = e[2]
        arr[i].b = e[3] if len(e) > 3 else 0
    return arr, len(ev)


class RefInstance:
    """One reference organ (struct b_instance, src/global_inst.h:32-40)."""

    def __init__(self, freqs=None, rate=48000.0, seed=1, cfg="", target_ratio=None):
        L = lib()
        f = None
        if freqs is not None:
            f = (C.c_double * 128)(*[float(x) for x in freqs])
        else:
            f = (C.c_double * 128)(*tet12())
        tr = None
        if target_ratio is not None:
            tr = (C.c_double * 9)(*[float(x) for x in target_ratio])
        self._h = L.ref_create(rate, f, tr, seed, cfg.encode())
        if not self._h:
            raise MemoryError("ref_create failed")
        self._tt = None

    def close(self):
        if self._h:
            if self._tt is not None:
                lib().ref_free_tonegen_tables(C.byref(self._tt))
                self._tt = None
            lib().ref_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def event(self, type_, a=0, b=0):
        if lib().ref_event(self._h, type_, a, b) != 0:
            raise ValueError("bad event type %r" % (type_,))

    def render(self, n_fragments, flags=0, events=()):
        """-> (L, R) float32 arrays of n_fragments*128 samples."""
        outl = np.zeros(n_fragments * 128, dtype=np.float32)
        outr = np.zeros(n_fragments * 128, dtype=np.float32)
        arr, n = events_array(events)
        lib().ref_render(self._h, n_fragments, flags, arr, n,
                         outl.ctypes.data_as(C.c_void_p), outr.ctypes.data_as(C.c_void_p))
        return outl, outr

    def render_midi(self, n_fragments, flags=0, messages=()):
        """Render with raw MIDI: messages = [(fragment, bytes)] sorted by fragment, each fed
        to the reference's parse_raw_midi_data() before its fragment.  -> (L, R)."""
        outl = np.zeros(n_fragments * 128, dtype=np.float32)
        outr = np.zeros(n_fragments * 128, dtype=np.float32)
        flat = []
        for fragment, raw in messages:
            raw = bytes(raw)
            flat += [fragment, len(raw)] + list(raw[:3]) + [0] * (3 - min(3, len(raw)))
        arr = (C.c_uint * max(1, len(flat)))(*flat)
        lib().ref_render_midi(self._h, n_fragments, flags, arr, len(flat) // 5,
                              outl.ctypes.data_as(C.c_void_p), outr.ctypes.data_as(C.c_void_p))
        return outl, outr

    def whirl_only(self, n_fragments, flags, mono, events=()):
        """whirlProc / whirlProc3 on the caller's mono buffer; only rotor events apply."""
        mono = np.ascontiguousarray(mono, dtype=np.float32)
        outl = np.zeros(n_fragments * 128, dtype=np.float32)
        outr = np.zeros(n_fragments * 128, dtype=np.float32)
        arr, n = events_array(events)
        lib().ref_whirl_only(self._h, n_fragments, flags, mono.ctypes.data_as(C.c_void_p), arr, n,
                             outl.ctypes.data_as(C.c_void_p), outr.ctypes.data_as(C.c_void_p))
        return outl, outr

    def whirl2(self, n_fragments, mono):
        """whirlProc2 on the caller's mono buffer -> (L, R, [HL, HR, DL, DR])."""
        mono = np.ascontiguousarray(mono, dtype=np.float32)
        outs = [np.zeros(n_fragments * 128, dtype=np.float32) for _ in range(6)]
        lib().ref_whirl2(self._h, n_fragments, mono.ctypes.data_as(C.c_void_p),
                         *[o.ctypes.data_as(C.c_void_p) for o in outs])
        return outs[0], outs[1], outs[2:]

    def load_programmes(self, text):
        """Programme definitions (.pgm syntax) through the reference's own parser."""
        if lib().ref_load_programmes(self._h, text.encode()) != 0:
            raise ValueError("the reference rejected the programme text")

    def midi_key(self, channel, note):
        """Engine key the reference maps (channel, note) to, -1 if none (observed through
        activeKeys[], the key tables being private to midi.cpp)."""
        return lib().ref_midi_key(self._h, channel, note)

    def active_wheels(self):
        return lib().ref_active_wheels(self._h)

    def tonegen_tables(self):
        """struct b_tonegen flattened into the boundary's table format (kept alive by self)."""
        if self._tt is None:
            self._tt = cabi.TonegenTables()
            lib().ref_dump_tonegen_tables(self._h, C.byref(self._tt))
        return self._tt

    def whirl_tables(self):
        wt = cabi.WhirlTables()
        lib().ref_dump_whirl_tables(self._h, C.byref(wt))
        return wt

    def export_state(self):
        st = cabi.StateExport()
        lib().ref_export_state(self._h, C.byref(st))
        return st
