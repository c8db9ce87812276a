"""Host-side instance specification and table building (include/tbf_tables.h).

`InstanceSpec` is the resolved form of the reference's inputs: the 128 MTS-ESP
note frequencies, cfg text (osc.*, scanner.*, whirl.*), the rand() seed and the
sample rate.  `build()` runs the C++ table builder in libtbf_gpu.so -- the
product's own restatement of initToneGenerator / init_vibrato / initWhirl.
"""
import ctypes as C

import numpy as np

from . import cabi
from .engine import lib as _gpu_lib

_bound = False


def _lib():
    global _bound
    L = _gpu_lib()
    if not _bound:
        L.tbf_tonegen_spec_default.argtypes = [C.POINTER(cabi.TonegenSpec), C.c_double]
        L.tbf_tonegen_spec_default.restype = None
        L.tbf_whirl_spec_default.argtypes = [C.POINTER(cabi.WhirlSpec), C.c_double]
        L.tbf_whirl_spec_default.restype = None
        L.tbf_freqs_12tet.argtypes = [C.POINTER(C.c_double)]
        L.tbf_freqs_12tet.restype = None
        L.tbf_freqs_from_scl.argtypes = [C.c_char_p, C.POINTER(C.c_double)]
        L.tbf_freqs_from_ratios.argtypes = [C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double)]
        L.tbf_scl_ratios.argtypes = [C.c_char_p, C.POINTER(C.c_double), C.c_int]
        L.tbf_spec_apply_cfg.argtypes = [C.POINTER(cabi.TonegenSpec), C.POINTER(cabi.WhirlSpec),
                                         C.c_char_p, C.c_char_p]
        L.tbf_build_tonegen_tables.argtypes = [C.POINTER(cabi.TonegenSpec),
                                               C.POINTER(cabi.TonegenTables)]
        L.tbf_build_tonegen_tables_synth.argtypes = [C.POINTER(cabi.TonegenSpec), C.POINTER(cabi.TonegenTables),
                                                     C.POINTER(cabi.WaveSynth)]
        L.tbf_wave_sample.argtypes = [C.POINTER(cabi.WaveSynth), C.c_int, C.c_uint32, C.c_int]
        L.tbf_wave_sample.restype = C.c_float
        L.tbf_dither_bits.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p]
        L.tbf_dither_bits.restype = None
        L.tbf_free_tonegen_tables.argtypes = [C.POINTER(cabi.TonegenTables)]
        L.tbf_free_tonegen_tables.restype = None
        L.tbf_build_whirl_tables.argtypes = [C.POINTER(cabi.WhirlSpec), C.POINTER(cabi.WhirlTables)]
        L.tbf_free_whirl_tables.argtypes = [C.POINTER(cabi.WhirlTables)]
        L.tbf_free_whirl_tables.restype = None
        L.tbf_extend_frequencies.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double),
                                             C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.tbf_iir_coefficients.argtypes = [C.POINTER(C.c_float), C.c_int] + [C.c_double] * 4
        L.tbf_srand.argtypes = [C.c_void_p, C.c_uint32]
        L.tbf_srand.restype = None
        L.tbf_rand.argtypes = [C.c_void_p]
        L.tbf_rand.restype = C.c_int32
        _bound = True
    return L


def freqs_12tet():
    f = (C.c_double * 128)()
    _lib().tbf_freqs_12tet(f)
    return np.array(f, dtype=np.float64)


def freqs_from_scl(text):
    """Scala .scl text -> 128 note frequencies (degree 0 at MIDI 60 = 261.6255653 Hz)."""
    f = (C.c_double * 128)()
    rc = _lib().tbf_freqs_from_scl(text.encode("latin-1", "replace"), f)
    if rc != cabi.OK:
        raise ValueError("unparsable .scl (%d)" % rc)
    return np.array(f, dtype=np.float64)


def scl_ratios(text):
    """Scala .scl text -> the scale's degrees as frequency ratios (last = period)."""
    buf = (C.c_double * 4096)()
    n = _lib().tbf_scl_ratios(text.encode("latin-1", "replace"), buf, 4096)
    if n < 1:
        raise ValueError("unparsable .scl (%d)" % n)
    return np.array(buf[:n], dtype=np.float64)


def freqs_from_ratios(degrees):
    d = np.ascontiguousarray(degrees, dtype=np.float64)
    f = (C.c_double * 128)()
    rc = _lib().tbf_freqs_from_ratios(d.ctypes.data_as(C.POINTER(C.c_double)), len(d), f)
    if rc != cabi.OK:
        raise ValueError("bad scale degrees (%d)" % rc)
    return np.array(f, dtype=np.float64)


def extend_frequencies(freqs128):
    f = (C.c_double * 128)(*[float(x) for x in freqs128])
    out = (C.c_double * cabi.NOF_FREQS)()
    ss = C.c_int(0)
    per = C.c_float(0)
    _lib().tbf_extend_frequencies(f, out, C.byref(ss), C.byref(per))
    return np.array(out, dtype=np.float64), ss.value, per.value


def iir_coefficients(type_, hz, q, gain_db, rate=48000.0):
    """setIIRFilter / eqCompute (src/whirl.cpp:147-172): [a1, a2, b0, b1, b2] as float32, or
    None when the reference would keep its previous coefficients (parameter out of range)."""
    c5 = (C.c_float * 5)()
    rc = _lib().tbf_iir_coefficients(c5, int(type_), hz, q, gain_db, rate)
    if rc < 0:
        raise ValueError("tbf_iir_coefficients failed (%d)" % rc)
    return list(c5) if rc == 1 else None


def parse_cfg_lines(text):
    """cfg text -> [(key, value)] the way the reference tokenises a line
    (parseConfigurationLine, src/cfgParser.cpp:94-160): '#' comments, key=value,
    surrounding blanks stripped."""
    out = []
    for line in text.splitlines():
        s = line.strip()
        if not s or s.startswith("#") or s.startswith("="):
            continue
        key, _, val = s.partition("=")
        val = val.split("#", 1)[0]
        out.append((key.strip(), val.strip()))
    return out


class InstanceSpec:
    def __init__(self, freqs=None, rate=48000.0, seed=1, cfg="", target_ratio=None):
        L = _lib()
        self.tg = cabi.TonegenSpec()
        self.wh = cabi.WhirlSpec()
        L.tbf_tonegen_spec_default(C.byref(self.tg), rate)
        L.tbf_whirl_spec_default(C.byref(self.wh), rate)
        if freqs is not None:
            for i in range(128):
                self.tg.freqs[i] = float(freqs[i])
        if target_ratio is not None:
            for i in range(9):
                self.tg.target_ratio[i] = float(target_ratio[i])
        self.tg.seed = seed
        self.rejected = []   # out-of-range values: ignored, like the reference does
        self.unclaimed = []  # keys of other modules (midi.*, reverb.*, ...): not this path
        if cfg:
            self.apply_cfg(cfg)

    def apply_cfg(self, text):
        L = _lib()
        for key, val in parse_cfg_lines(text):
            rc = L.tbf_spec_apply_cfg(C.byref(self.tg), C.byref(self.wh), key.encode(), val.encode())
            if rc == 0:
                self.unclaimed.append((key, val))
            elif rc < 0:
                self.rejected.append((key, val))

    def build_tonegen(self):
        """-> TonegenTables only (free with free_tables(tt, None))."""
        tt = cabi.TonegenTables()
        rc = _lib().tbf_build_tonegen_tables(C.byref(self.tg), C.byref(tt))
        if rc != cabi.OK:
            raise ValueError("tbf_build_tonegen_tables failed (%d)" % rc)
        return tt

    def build_tonegen_synth(self):
        """-> (TonegenTables without wave samples, WaveSynth): everything initToneGenerator builds
        except the wavetables, which tbf_gpu_tonegen_tables_create_synth makes on the device."""
        tt = cabi.TonegenTables()
        ws = cabi.WaveSynth()
        rc = _lib().tbf_build_tonegen_tables_synth(C.byref(self.tg), C.byref(tt), C.byref(ws))
        if rc != cabi.OK:
            raise ValueError("tbf_build_tonegen_tables_synth failed (%d)" % rc)
        return tt, ws

    def build(self):
        """-> (TonegenTables, WhirlTables); free with free_tables()."""
        L = _lib()
        tt = cabi.TonegenTables()
        rc = L.tbf_build_tonegen_tables(C.byref(self.tg), C.byref(tt))
        if rc != cabi.OK:
            raise ValueError("tbf_build_tonegen_tables failed (%d)" % rc)
        wt = cabi.WhirlTables()
        rc = L.tbf_build_whirl_tables(C.byref(self.wh), C.byref(wt))
        if rc != cabi.OK:
            L.tbf_free_tonegen_tables(C.byref(tt))
            raise ValueError("tbf_build_whirl_tables failed (%d)" % rc)
        return tt, wt


def free_tables(tt, wt):
    L = _lib()
    if tt is not None:
        L.tbf_free_tonegen_tables(C.byref(tt))
    if wt is not None:
        L.tbf_free_whirl_tables(C.byref(wt))


def table_digest(tt):
    """blake2b-128 over every table of a tbf_tonegen_tables (value fields and arrays)."""
    import hashlib
    h = hashlib.blake2b(digest_size=16)
    for fname, _ in tt._fields_:
        v = getattr(tt, fname)
        if isinstance(v, C._Pointer):
            continue
        if isinstance(v, C.Array):
            h.update(bytes(v))
        elif isinstance(v, float):
            h.update(np.float64(v).tobytes())
        else:
            h.update(np.int64(v).tobytes())
    nc = tt.contrib_start[cabi.MAX_KEYS]
    for ptr, n, dt in ((tt.wave_data, tt.wave_total, np.uint32), (tt.contrib_wheel, nc, np.uint16),
                       (tt.contrib_bus, nc, np.uint16), (tt.contrib_level, nc, np.uint32)):
        h.update(np.ctypeslib.as_array(ptr, shape=(int(n),)).view(dt).tobytes())
    d = h.digest()
    return int.from_bytes(d[:8], "little"), int.from_bytes(d[8:], "little")
