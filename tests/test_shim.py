"""libtbf_shim.so: the reference's per-instance engine API (include/tbf_shim.h) over the CUDA
path.  The GPU test is the hosts' synthSound() loop (src/clap.cpp:244-270): setter calls between
fragments, oscGenerateFragment -> (preamp / reverb place) -> whirlProc3, one fragment at a time,
compared bit for bit with the reference engine driven by the same calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tunebfree_b200 import cabi, workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tunebfree_b200", "libtbf_shim.so")


class ConfigContext(C.Structure):
    _fields_ = [("fname", C.c_char_p), ("linenr", C.c_int), ("name", C.c_char_p), ("value", C.c_char_p)]


def _declared():
    txt = open(os.path.join(ROOT, "include", "tbf_shim.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", txt)))


def test_shim_exports_the_reference_names():
    L = C.CDLL(LIB)
    names = _declared()
    for must in ("allocTonegen", "initToneGenerator", "oscKeyOn", "oscKeyOff", "setDrawBar",
                 "oscGenerateFragment", "allocWhirl", "initWhirl", "useRevOption", "whirlProc",
                 "whirlProc3", "setVibratoFromInt", "oscConfig", "whirlConfig"):
        assert must in names
    for n in names:
        assert hasattr(L, n), "libtbf_shim.so does not export %s" % n


def test_shim_config_return_convention():
    """oscConfig / scannerConfig / whirlConfig: 1 = claimed, 0 = not this module's key, -1 = bad
    value -- no GPU needed until init."""
    L = C.CDLL(LIB)
    L.allocTonegen.restype = C.c_void_p
    L.allocWhirl.restype = C.c_void_p
    t, w = L.allocTonegen(), L.allocWhirl()

    def cc(k, v):
        return C.byref(ConfigContext(b"<test>", 1, k, v))
    assert L.oscConfig(C.c_void_p(t), cc(b"osc.perc.fast", b"0.5")) == 1
    assert L.oscConfig(C.c_void_p(t), cc(b"whirl.horn.level", b"0.5")) == 0
    assert L.scannerConfig(C.c_void_p(t), cc(b"scanner.hz", b"6.9")) == 1
    assert L.whirlConfig(C.c_void_p(w), cc(b"whirl.horn.level", b"0.6")) == 1
    assert L.whirlConfig(C.c_void_p(w), cc(b"osc.perc.fast", b"0.5")) == 0
    assert L.whirlConfig(C.c_void_p(w), cc(b"whirl.speed-preset", b"fast")) == -1
    L.freeToneGenerator(C.c_void_p(t))
    L.freeWhirl(C.c_void_p(w))


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", ["", "osc.perc.fast=0.6\nosc.attack.click.level=0.7\nscanner.hz=6.2\n"
                                     "whirl.horn.level=0.6\nwhirl.drum.radius=20\n"])
def test_synthsound_loop_bit_exact(ref, cfg):
    L = C.CDLL(LIB)
    for f in ("allocTonegen", "allocWhirl"):
        getattr(L, f).restype = C.c_void_p
    L.tbf_shim_last_error.restype = C.c_char_p
    L.initToneGenerator.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
    L.initWhirl.argtypes = [C.c_void_p, C.c_void_p, C.c_double]
    L.oscGenerateFragment.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    L.whirlProc3.argtypes = [C.c_void_p] * 6 + [C.c_size_t]
    L.whirlProc.argtypes = [C.c_void_p] * 4 + [C.c_size_t]
    L.oscKeyOn.argtypes = L.oscKeyOff.argtypes = [C.c_void_p, C.c_short, C.c_short]
    L.setDrawBar.argtypes = [C.c_void_p, C.c_int, C.c_uint]
    for f in ("setPercussionEnabled", "setPercussionVolume", "setPercussionFast", "setPercussionFirst",
              "setVibratoUpper", "setVibratoLower", "setVibratoFromInt"):
        getattr(L, f).argtypes = [C.c_void_p, C.c_int]
    L.useRevOption.argtypes = [C.c_void_p, C.c_int, C.c_int]

    n_frag = 150
    ev = workloads.instance_events(3, n_frag, config=3, restrike=60, hold=40)
    ev += [(90, cabi.EV_REV_OPTION, 4, 0), (100, cabi.EV_DRAWBAR, 2, 5), (110, cabi.EV_VIB_KNOB, 2, 0)]
    ev.sort(key=lambda e: e[0])
    freqs = ref.tet12() * 1.003
    r = ref.RefInstance(freqs=freqs, seed=5, cfg=cfg)
    rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, ev)
    r.close()

    # allocSynth / initSynth order of the hosts (b_synth/lv2.cpp:336-353, :164-193)
    t, w = L.allocTonegen(), L.allocWhirl()
    for ln, line in enumerate(cfg.splitlines()):
        k, v = line.split("=")
        cc = ConfigContext(b"<test>", ln + 1, k.encode(), v.encode())
        claimed = (L.oscConfig(C.c_void_p(t), C.byref(cc)) + L.scannerConfig(C.c_void_p(t), C.byref(cc))
                   + L.whirlConfig(C.c_void_p(w), C.byref(cc)))
        assert claimed == 1, line
    L.tbf_shim_set_tuning(C.c_void_p(t), (C.c_double * 128)(*freqs))
    L.tbf_shim_set_seed(C.c_void_p(t), 5)
    L.initToneGenerator(t, None, 48000.0, None)
    L.initVibrato(C.c_void_p(t), None)
    L.initWhirl(w, None, 48000.0)
    assert L.tbf_shim_status() == 0, L.tbf_shim_last_error()

    class BInstance(C.Structure):  # src/global_inst.h:32-40
        _fields_ = [(n, C.c_void_p) for n in ("reverb", "whirl", "synth", "progs", "midicfg", "preamp", "state")]
    inst = BInstance(None, w, t, None, None, None, None)
    upper = [None] * 9

    def set_drawbar(bus, setting):
        """the nine fragment-0 upper drawbars go through setDrawBars(inst, 0, ...), later
        single changes through setDrawBar"""
        if upper is not None and bus < 9 and upper[bus] is None:
            upper[bus] = setting
            if all(u is not None for u in upper):
                L.setDrawBars(C.byref(inst), 0, (C.c_uint * 9)(*upper))
            return
        L.setDrawBar(t, bus, setting)

    call = {
        cabi.EV_KEY_ON: lambda a, b: L.oscKeyOn(t, a, a),
        cabi.EV_KEY_OFF: lambda a, b: L.oscKeyOff(t, a, a),
        cabi.EV_DRAWBAR: set_drawbar,
        cabi.EV_PERC_ENABLE: lambda a, b: L.setPercussionEnabled(t, a),
        cabi.EV_PERC_SOFT: lambda a, b: L.setPercussionVolume(t, a),
        cabi.EV_PERC_FAST: lambda a, b: L.setPercussionFast(t, a),
        cabi.EV_PERC_FIRST: lambda a, b: L.setPercussionFirst(t, a),
        cabi.EV_VIB_UPPER: lambda a, b: L.setVibratoUpper(t, a),
        cabi.EV_VIB_KNOB: lambda a, b: L.setVibratoFromInt(t, a),
        cabi.EV_REV_OPTION: lambda a, b: L.useRevOption(w, a, 2),
    }
    buf_a = np.zeros(128, dtype=np.float32)
    tmp = np.zeros((2, 128), dtype=np.float32)
    out = np.zeros((2, n_frag * 128), dtype=np.float32)
    k = 0
    for f in range(n_frag):
        while k < len(ev) and ev[k][0] <= f:
            call[ev[k][1]](ev[k][2], ev[k][3] if len(ev[k]) > 3 else 0)
            k += 1
        L.oscGenerateFragment(t, buf_a.ctypes.data, 128)
        # preamp (clean) and reverb (dry) would run here on the host's bufA
        L.whirlProc3(w, buf_a.ctypes.data, out[0, f * 128:].ctypes.data, out[1, f * 128:].ctypes.data,
                     tmp[0].ctypes.data, tmp[1].ctypes.data, 128)
    assert L.tbf_shim_status() == 0, L.tbf_shim_last_error()
    L.freeToneGenerator(C.c_void_p(t))
    L.freeWhirl(C.c_void_p(w))
    L.tbf_shim_shutdown()
    assert float(np.abs(rl).max()) > 0.05
    assert np.array_equal(out[0].view(np.uint32), rl.view(np.uint32))
    assert np.array_equal(out[1].view(np.uint32), rr.view(np.uint32))
