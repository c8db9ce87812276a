"""Host-side mirror of the reference's engine API over the C-ABI library.

The reference's operator API is a set of C-style functions over heap structs
(src/tonegen.h:565-597, src/vibrato.h:70-79, src/whirl.h:223-272).  This module
keeps those names (``oscKeyOn``, ``setDrawBar``, ``useRevOption`` ...) as
methods of :class:`Organ`; they record control events that the batched GPU
renderer applies at fragment boundaries, exactly where every reference host
applies them (b_synth/lv2.cpp:1130-1136, src/clap.cpp:251).

There is NO CPU fallback: if ``libtbf_gpu.so`` is missing or no CUDA device is
usable, construction raises.
"""
import ctypes as C
import os

import numpy as np

from . import cabi

_HERE = os.path.dirname(os.path.abspath(__file__))
# TBF_LIB: another build of the same library (kernel experiments); the product is libtbf_gpu.so
LIB_PATH = os.environ.get("TBF_LIB") or os.path.join(_HERE, "libtbf_gpu.so")
_lib = None


class TbfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("tbf_gpu error %d: %s" % (code, msg))
        self.code = code


def lib():
    """Load the C-ABI library (built in-tree by ``__graft_entry__.build()``)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.tbf_gpu_create.argtypes = [C.POINTER(vp), C.c_int]
        L.tbf_gpu_destroy.argtypes = [vp]
        L.tbf_gpu_last_error.argtypes = [vp]
        L.tbf_gpu_last_error.restype = C.c_char_p
        L.tbf_gpu_tonegen_tables_create.argtypes = [vp, C.POINTER(cabi.TonegenTables),
                                                    C.POINTER(C.c_int)]
        L.tbf_gpu_whirl_tables_create.argtypes = [vp, C.POINTER(cabi.WhirlTables),
                                                  C.POINTER(C.c_int)]
        L.tbf_gpu_whirl_tables_info.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                                C.POINTER(C.c_int)]
        L.tbf_gpu_instance_create.argtypes = [vp, C.c_int, C.c_int, C.c_uint32,
                                              C.POINTER(C.c_uint32)]
        L.tbf_gpu_instance_count.argtypes = [vp]
        L.tbf_gpu_instance_reset.argtypes = [vp, C.c_uint32, C.c_uint32]
        L.tbf_gpu_render_batch.argtypes = [vp, C.POINTER(cabi.BatchDesc)]
        L.tbf_gpu_render_batch_launch.argtypes = [vp, C.POINTER(cabi.BatchDesc)]
        L.tbf_gpu_synchronize.argtypes = [vp]
        L.tbf_gpu_batch_prepare.argtypes = [vp, C.POINTER(cabi.BatchDesc)]
        L.tbf_gpu_batch_launch.argtypes = [vp]
        L.tbf_gpu_host_alloc.argtypes = [vp, C.c_size_t]
        L.tbf_gpu_host_alloc.restype = C.c_void_p
        L.tbf_gpu_host_free.argtypes = [vp, C.c_void_p]
        L.tbf_gpu_host_free.restype = None
        L.tbf_gpu_set_host_chunk.argtypes = [vp, C.c_uint32]
        L.tbf_gpu_timer_start.argtypes = [vp]
        L.tbf_gpu_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
        L.tbf_gpu_last_launch_ms.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.tbf_gpu_export_state.argtypes = [vp, C.c_uint32, C.POINTER(cabi.StateExport)]
        L.tbf_gpu_set_stage_profiling.argtypes = [vp, C.c_int]
        L.tbf_gpu_stage_times.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.tbf_gpu_tonegen_tables_create_synth.argtypes = [vp, C.POINTER(cabi.TonegenTables),
                                                          C.POINTER(cabi.WaveSynth), C.POINTER(C.c_int)]
        L.tbf_gpu_synth_stats.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.tbf_gpu_tonegen_tables_read_waves.argtypes = [vp, C.c_int, C.c_void_p, C.c_uint64]
        L.tbf_gpu_organs_create.argtypes = [vp, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                            C.POINTER(C.c_int), C.POINTER(C.c_uint32), C.POINTER(C.c_double)]
        L.tbf_gpu_render_small_noalloc.argtypes = [vp, C.POINTER(cabi.BatchDesc)]
        L.tbf_gpu_noalloc_scratch_bytes.restype = C.c_ulonglong
        L.tbf_gpu_instance_destroy.argtypes = [vp, C.c_uint32, C.c_uint32]
        L.tbf_gpu_tonegen_tables_destroy.argtypes = [vp, C.c_int]
        L.tbf_gpu_whirl_tables_destroy.argtypes = [vp, C.c_int]
        L.tbf_gpu_resource_counts.argtypes = [vp] + [C.POINTER(C.c_uint32)] * 4
        L.tbf_gpu_instance_stats.argtypes = [vp, C.POINTER(C.c_uint32), C.c_uint32,
                                             C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.tbf_gpu_get_host_chunk.argtypes = [vp]
        L.tbf_gpu_get_host_chunk.restype = C.c_uint32
        L.tbf_gpu_host_alloc2.argtypes = [vp, C.c_size_t, C.c_uint]
        L.tbf_gpu_host_alloc2.restype = C.c_void_p
        L.tbf_gpu_measure_d2h.argtypes = [vp, C.c_size_t, C.c_int, C.POINTER(C.c_double)]
        L.tbf_gpu_event_index.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                          C.POINTER(C.c_uint64), C.POINTER(C.c_char_p)]
        _lib = L
    return _lib


def events_to_array(per_slot_events):
    """per_slot_events: list (one entry per batch slot) of [(fragment, type, a, b)].
    Returns (ctypes array of tbf_event sorted by (slot, fragment), count).
    The order of events inside one (slot, fragment) is preserved."""
    n = sum(len(e) for e in per_slot_events)
    arr = (cabi.Event * max(1, n))()
    i = 0
    for slot, evs in enumerate(per_slot_events):
        for e in sorted(evs, key=lambda t: t[0]):  # stable
            arr[i].instance = slot
            arr[i].fragment = e[0]
            arr[i].type = e[1]
            arr[i].a = e[2]
            arr[i].b = e[3] if len(e) > 3 else 0
            i += 1
    return arr, n


def events_from_numpy(ev):
    """ev: structured/2-D int array [n, 5] = (slot, fragment, type, a, b), already
    sorted by (slot, fragment) -> ctypes array sharing no memory with ev."""
    ev = np.ascontiguousarray(ev, dtype=np.int32)
    n = ev.shape[0]
    arr = (cabi.Event * max(1, n))()
    if n:
        C.memmove(arr, ev.ctypes.data, n * C.sizeof(cabi.Event))
    return arr, n


def event_index(ev, n_instances, n_fragments):
    """tbf_gpu_event_index on an [n, 5] int32 array: the per-slot start indices (n_instances + 1), or
    TbfError naming the rule the list breaks.  Needs no GPU."""
    ev = np.ascontiguousarray(ev, dtype=np.int32).reshape(-1, 5)
    start = np.zeros(n_instances + 1, dtype=np.uint64)
    why = C.c_char_p()
    rc = lib().tbf_gpu_event_index(ev.ctypes.data, ev.shape[0], n_instances, n_fragments,
                                   start.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(why))
    if rc != 0:
        raise TbfError(rc, (why.value or b"").decode())
    return start


def events_from_numpy_pinned(ctx, ev):
    """As events_from_numpy, the array in page-locked memory of `ctx` (a plain DMA instead of a staged
    copy when the batch is prepared) -> ((array, n), address for ctx.host_free)."""
    ev = np.ascontiguousarray(ev, dtype=np.int32)
    n = ev.shape[0]
    host, addr = ctx.host_alloc((max(1, n) * C.sizeof(cabi.Event),), dtype=np.uint8)
    if n:
        C.memmove(addr, ev.ctypes.data, n * C.sizeof(cabi.Event))
    arr = (cabi.Event * max(1, n)).from_address(addr)
    return (arr, n), addr


def checksum_host(out_l, out_r):
    """The kernel's position-keyed additive hash of one instance's L/R bit patterns."""
    def mix(x):
        x = (x + np.uint64(0x9E3779B97F4A7C15))
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))
    with np.errstate(over="ignore"):
        idx = np.arange(out_l.shape[0], dtype=np.uint64) << np.uint64(32)
        a = mix(idx | out_l.view(np.uint32).astype(np.uint64))
        b = mix((idx | out_r.view(np.uint32).astype(np.uint64)) ^ np.uint64(0x5555555555555555))
        return int((a.sum(dtype=np.uint64) + b.sum(dtype=np.uint64)) & np.uint64(0xFFFFFFFFFFFFFFFF))


class Context:
    """tbf_ctx: one GPU."""

    def __init__(self, device=0):
        L = lib()
        h = C.c_void_p()
        rc = L.tbf_gpu_create(C.byref(h), device)
        if rc != cabi.OK:
            raise TbfError(rc, "tbf_gpu_create failed (no usable CUDA device %d?)" % device)
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            lib().tbf_gpu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != cabi.OK:
            raise TbfError(rc, (lib().tbf_gpu_last_error(self._h) or b"").decode())

    def tonegen_tables(self, tables):
        h = C.c_int(-1)
        self._ck(lib().tbf_gpu_tonegen_tables_create(self._h, C.byref(tables), C.byref(h)))
        return h.value

    def tonegen_tables_synth(self, tables, synth):
        """Upload a table set whose wavetables the device synthesises (tables.wave_data NULL)."""
        h = C.c_int(-1)
        self._ck(lib().tbf_gpu_tonegen_tables_create_synth(self._h, C.byref(tables), C.byref(synth),
                                                           C.byref(h)))
        return h.value

    def synth_stats(self):
        """-> (samples flagged as rounding-ambiguous, of those: differed from the host's libm)"""
        a, b = C.c_uint64(0), C.c_uint64(0)
        self._ck(lib().tbf_gpu_synth_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def read_waves(self, handle, n_floats):
        out = np.zeros(int(n_floats), dtype=np.float32)
        self._ck(lib().tbf_gpu_tonegen_tables_read_waves(self._h, handle, out.ctypes.data, int(n_floats)))
        return out

    def create_organs(self, specs, whirl_handle, threads=0, device_waves=True):
        """tbf_gpu_organs_create: one organ per cabi.TonegenSpec of `specs` (a ctypes array), built
        on `threads` host threads, uploaded and created in order.
        -> (instance ids, tonegen handles, (wall s, waiting-for-builders s, upload s))"""
        n = len(specs)
        ids = (C.c_uint32 * n)()
        tgh = (C.c_int * n)()
        st = (C.c_double * 3)()
        self._ck(lib().tbf_gpu_organs_create(self._h, C.cast(specs, C.c_void_p), n, whirl_handle, threads,
                                             1 if device_waves else 0, tgh, ids, st))
        return np.array(ids, dtype=np.uint32), list(tgh), tuple(st)

    def whirl_tables(self, tables):
        h = C.c_int(-1)
        self._ck(lib().tbf_gpu_whirl_tables_create(self._h, C.byref(tables), C.byref(h)))
        return h.value

    def whirl_tables_info(self, handle):
        """-> (gather_ok, ring_size, window)"""
        g, r, w = C.c_int(0), C.c_int(0), C.c_int(0)
        self._ck(lib().tbf_gpu_whirl_tables_info(self._h, handle, C.byref(g), C.byref(r), C.byref(w)))
        return g.value, r.value, w.value

    def create_instances(self, tg_handle, wh_handle, count=1):
        first = C.c_uint32(0)
        self._ck(lib().tbf_gpu_instance_create(self._h, tg_handle, wh_handle, count,
                                               C.byref(first)))
        return first.value

    def reset(self, first, count):
        self._ck(lib().tbf_gpu_instance_reset(self._h, first, count))

    def destroy_instances(self, first, count=1):
        self._ck(lib().tbf_gpu_instance_destroy(self._h, first, count))

    def destroy_tonegen_tables(self, handle):
        self._ck(lib().tbf_gpu_tonegen_tables_destroy(self._h, handle))

    def destroy_whirl_tables(self, handle):
        self._ck(lib().tbf_gpu_whirl_tables_destroy(self._h, handle))

    def resource_counts(self):
        """-> (live instances, instance slots, live tonegen sets, live whirl sets)"""
        v = [C.c_uint32(0) for _ in range(4)]
        self._ck(lib().tbf_gpu_resource_counts(self._h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def instance_stats(self, ids=None, n=None):
        """-> (fragments rendered, sum of active-list lengths) over the listed instances."""
        f, w = C.c_uint64(0), C.c_uint64(0)
        ida = None
        if ids is not None:
            n = len(ids)
            ida = (C.c_uint32 * n)(*[int(i) for i in ids])
        self._ck(lib().tbf_gpu_instance_stats(self._h, ida, n, C.byref(f), C.byref(w)))
        return f.value, w.value

    def measure_d2h(self, nbytes=1 << 30, reps=3):
        """Plain contiguous device -> page-locked host copy rate of this GPU's link, GB/s."""
        g = C.c_double(0)
        self._ck(lib().tbf_gpu_measure_d2h(self._h, nbytes, reps, C.byref(g)))
        return g.value

    def host_chunk(self):
        return lib().tbf_gpu_get_host_chunk(self._h)

    def instance_count(self):
        return lib().tbf_gpu_instance_count(self._h)

    def _desc(self, n_instances, n_fragments, events, flags, ids, out_l, out_r, stride, checksums):
        d = cabi.BatchDesc()
        d.n_instances = n_instances
        d.n_fragments = n_fragments
        keep = []
        if ids is not None:
            ida = (C.c_uint32 * n_instances).from_buffer_copy(
                np.ascontiguousarray(np.asarray(ids, dtype=np.uint32)[:n_instances]))
            keep.append(ida)
            d.instance_ids = ida
        if isinstance(events, tuple):
            arr, n = events
        else:
            arr, n = events_to_array(events if events is not None else [])
        keep.append(arr)
        d.events = arr
        d.n_events = n
        d.out_l = out_l
        d.out_r = out_r
        d.out_stride = stride
        d.flags = flags
        if checksums is not None:
            d.checksums = checksums.ctypes.data_as(C.POINTER(C.c_uint64))
        return d, keep

    def render(self, n_instances, n_fragments, events=None, flags=0, ids=None, checksums=False,
               inputs=None, stems=False):
        """Host-buffer render through tbf_gpu_render_batch.
        events: per-slot lists [(fragment, type, a, b)] or a prebuilt (array, n).
        inputs: float32 [n_instances, n_fragments*128] mono buffers -> Leslie only
        (TBF_RENDER_WHIRL_INPUT, whirlProc / whirlProc3 on the caller's audio).
        Returns (L, R[, checksums]) with L, R of shape [n_instances, n_fragments*128]."""
        frames = n_fragments * cabi.FRAGMENT
        out_l = np.empty((n_instances, frames), dtype=np.float32)
        out_r = np.empty((n_instances, frames), dtype=np.float32)
        cs = np.zeros(n_instances, dtype=np.uint64) if checksums else None
        if inputs is not None:
            inputs = np.ascontiguousarray(inputs, dtype=np.float32).reshape(n_instances, frames)
            flags |= cabi.RENDER_WHIRL_INPUT
        d, keep = self._desc(n_instances, n_fragments, events, flags, ids,
                             out_l.ctypes.data, out_r.ctypes.data, frames, cs)
        if inputs is not None:
            d.in_ = inputs.ctypes.data
            d.in_stride = frames
        st = None
        if stems:  # whirlProc2's outHL outHR outDL outDR
            st = np.empty((4, n_instances, frames), dtype=np.float32)
            d.flags |= cabi.RENDER_STEMS
            for j in range(4):
                d.stems[j] = st[j].ctypes.data
        self._ck(lib().tbf_gpu_render_batch(self._h, C.byref(d)))
        del keep
        res = (out_l, out_r)
        if checksums:
            res += (cs,)
        if stems:
            res += (st,)
        return res

    def render_device(self, n_instances, n_fragments, events, flags, out_l_ptr, out_r_ptr, stride,
                      ids=None, checksums=None, launch_only=False):
        """Device-buffer render (pointers are raw device addresses, e.g. torch data_ptr())."""
        flags |= cabi.RENDER_OUT_DEVICE if out_l_ptr else cabi.RENDER_NO_OUTPUT
        d, keep = self._desc(n_instances, n_fragments, events, flags, ids,
                             out_l_ptr, out_r_ptr, stride, checksums)
        fn = lib().tbf_gpu_render_batch_launch if launch_only else lib().tbf_gpu_render_batch
        self._ck(fn(self._h, C.byref(d)))
        self._keep = keep

    def render_small_noalloc(self, n_instances, n_fragments, events, flags, out_l_ptr, out_r_ptr, stride,
                             ids=None, checksums=None):
        """tbf_gpu_render_small_noalloc: a small batch with all scratch in the library's static
        device array (no allocation); outputs are raw device addresses."""
        d, keep = self._desc(n_instances, n_fragments, events, flags | cabi.RENDER_OUT_DEVICE, ids,
                             out_l_ptr, out_r_ptr, stride, checksums)
        self._ck(lib().tbf_gpu_render_small_noalloc(self._h, C.byref(d)))
        del keep

    def prepare(self, n_instances, n_fragments, events, flags, out_l_ptr=None, out_r_ptr=None,
                stride=0, ids=None, checksums=None):
        """tbf_gpu_batch_prepare: upload events once; launch() may then be repeated.
        Pointers are raw addresses (host, or device with RENDER_OUT_DEVICE)."""
        if not out_l_ptr:
            flags |= cabi.RENDER_NO_OUTPUT
        d, keep = self._desc(n_instances, n_fragments, events, flags, ids, out_l_ptr, out_r_ptr,
                             stride or n_fragments * cabi.FRAGMENT, checksums)
        self._ck(lib().tbf_gpu_batch_prepare(self._h, C.byref(d)))
        self._keep = (keep, checksums)

    def launch(self):
        self._ck(lib().tbf_gpu_batch_launch(self._h))

    def timer_start(self):
        self._ck(lib().tbf_gpu_timer_start(self._h))

    def timer_stop(self):
        ms = C.c_float(0)
        self._ck(lib().tbf_gpu_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def host_alloc(self, shape, dtype=np.float32, write_combined=False):
        """Page-locked host array (freed with host_free)."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = lib().tbf_gpu_host_alloc2(self._h, n, cabi.HOST_WRITE_COMBINED if write_combined else 0)
        if not p:
            raise MemoryError("tbf_gpu_host_alloc(%d) failed" % n)
        buf = (C.c_char * n).from_address(p)
        arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
        return arr, p

    def host_free(self, p):
        lib().tbf_gpu_host_free(self._h, p)

    def set_host_chunk(self, fragments):
        self._ck(lib().tbf_gpu_set_host_chunk(self._h, fragments))

    def synchronize(self):
        self._ck(lib().tbf_gpu_synchronize(self._h))

    def last_launch_ms(self):
        ms = C.c_float(0)
        n = C.c_int(0)
        self._ck(lib().tbf_gpu_last_launch_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def set_stage_profiling(self, on):
        """While on, every stage kernel runs alone between two CUDA events (no overlap)."""
        self._ck(lib().tbf_gpu_set_stage_profiling(self._h, 1 if on else 0))

    def stage_times(self):
        """-> {stage: (device ms, launches)} for stages T, R, W, D since profiling was set."""
        ms = (C.c_float * 4)()
        n = (C.c_int * 4)()
        self._ck(lib().tbf_gpu_stage_times(self._h, ms, n))
        return {k: (ms[i], n[i]) for i, k in enumerate("TRWD")}

    def export_state(self, instance):
        st = cabi.StateExport()
        self._ck(lib().tbf_gpu_export_state(self._h, instance, C.byref(st)))
        return st


class Organ:
    """Event recorder with the reference's function names for one batch slot
    (struct b_instance, src/global_inst.h:32-40).  ``fragment`` is the fragment
    index the following calls take effect at."""

    def __init__(self):
        self.events = []
        self.fragment = 0

    def at(self, fragment):
        self.fragment = fragment
        return self

    def _ev(self, t, a=0, b=0):
        self.events.append((self.fragment, t, int(a), int(b)))

    def oscKeyOn(self, key, realKey=None):
        self._ev(cabi.EV_KEY_ON, key)

    def oscKeyOff(self, key, realKey=None):
        self._ev(cabi.EV_KEY_OFF, key)

    def setDrawBar(self, bus, setting):
        self._ev(cabi.EV_DRAWBAR, bus, setting)

    def setDrawBars(self, manual, settings):
        for i, s in enumerate(settings):
            self.setDrawBar(manual * 9 + i, s)

    def setPercussionEnabled(self, on):
        self._ev(cabi.EV_PERC_ENABLE, 1 if on else 0)

    def setPercussionVolume(self, is_soft):
        self._ev(cabi.EV_PERC_SOFT, 1 if is_soft else 0)

    def setPercussionFast(self, is_fast):
        self._ev(cabi.EV_PERC_FAST, 1 if is_fast else 0)

    def setPercussionFirst(self, is_first):
        self._ev(cabi.EV_PERC_FIRST, 1 if is_first else 0)

    def setVibratoUpper(self, on):
        self._ev(cabi.EV_VIB_UPPER, 1 if on else 0)

    def setVibratoLower(self, on):
        self._ev(cabi.EV_VIB_LOWER, 1 if on else 0)

    def setVibratoFromInt(self, knob):
        self._ev(cabi.EV_VIB_KNOB, knob)

    def useRevOption(self, n):
        self._ev(cabi.EV_REV_OPTION, n)

    def setRevSelect(self, n):
        self._ev(cabi.EV_REV_SELECT, n)

    def setSwellPedal(self, u):
        self._ev(cabi.EV_SWELL, u)
