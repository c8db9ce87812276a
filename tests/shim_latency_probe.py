"""Developer probe (not a bench line): wall time of one host fragment through libtbf_shim.so
(oscGenerateFragment + whirlProc3 on one organ), the per-plugin-instance drop-in of INTEGRATION.md 2.
Usage: python tests/shim_latency_probe.py"""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = C.CDLL(os.path.join(ROOT, "tunebfree_b200", "libtbf_shim.so"))
L.allocTonegen.restype = C.c_void_p; L.allocWhirl.restype = C.c_void_p
L.initToneGenerator.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
L.initWhirl.argtypes = [C.c_void_p, C.c_void_p, C.c_double]
L.oscGenerateFragment.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
L.whirlProc3.argtypes = [C.c_void_p] * 6 + [C.c_size_t]
L.oscKeyOn.argtypes = [C.c_void_p, C.c_short, C.c_short]
L.setDrawBar.argtypes = [C.c_void_p, C.c_int, C.c_uint]
L.useRevOption.argtypes = [C.c_void_p, C.c_int, C.c_int]
t, w = L.allocTonegen(), L.allocWhirl()
L.initToneGenerator(t, None, 48000.0, None); L.initWhirl(w, None, 48000.0)
for b in range(9): L.setDrawBar(t, b, 8)
for k in (48, 52, 55, 60, 64, 67, 72): L.oscKeyOn(t, k, k)
L.useRevOption(w, 8, 2)
a = np.zeros(128, np.float32); o = np.zeros((4, 128), np.float32)
def frag():
    L.oscGenerateFragment(t, a.ctypes.data, 128)
    L.whirlProc3(w, a.ctypes.data, o[0].ctypes.data, o[1].ctypes.data, o[2].ctypes.data, o[3].ctypes.data, 128)
for _ in range(200): frag()
ts = []
for _ in range(2000):
    t0 = time.perf_counter(); frag(); ts.append(time.perf_counter() - t0)
ts = np.array(ts) * 1e6
print("shim oscGenerateFragment + whirlProc3 per 128-sample fragment: median %.0f us, p99 %.0f us, max %.0f us (budget 2667 us); peak %.3f; status %d"
      % (np.median(ts), np.percentile(ts, 99), ts.max(), float(np.abs(o[0]).max()), L.tbf_shim_status()))
