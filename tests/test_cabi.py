"""CPU-side checks of the C-ABI boundary: the library loads, exports every symbol
include/tbf_gpu.h declares, and the ctypes mirror matches the C struct sizes.
No device calls (no GPU here); the host-side event check of tbf_gpu_batch_prepare runs through
tbf_gpu_event_index."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tunebfree_b200 import cabi, engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    syms = set()
    for name in os.listdir(os.path.join(ROOT, "include")):
        if not name.endswith(".h") or name == "tbf_shim.h":  # libtbf_shim.so: tests/test_shim.py
            continue
        txt = open(os.path.join(ROOT, "include", name)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        for m in re.finditer(r"\b(tbf_[a-z0-9_]+)\s*\(", txt):
            syms.add(m.group(1))
    return syms


def test_library_exports_every_declared_symbol():
    L = engine.lib()
    syms = _declared_symbols()
    assert "tbf_gpu_render_batch" in syms and len(syms) >= 12
    for s in sorted(syms):
        assert hasattr(L, s), "libtbf_gpu.so does not export %s" % s


def test_struct_sizes_match_header():
    L = engine.lib()
    assert L.tbf_sizeof_tonegen_tables() == C.sizeof(cabi.TonegenTables)
    assert L.tbf_sizeof_whirl_tables() == C.sizeof(cabi.WhirlTables)
    assert L.tbf_sizeof_state_export() == C.sizeof(cabi.StateExport)
    assert L.tbf_sizeof_batch_desc() == C.sizeof(cabi.BatchDesc)
    assert L.tbf_sizeof_event() == C.sizeof(cabi.Event)
    from tunebfree_b200 import midi
    assert L.tbf_sizeof_midi_map() == C.sizeof(midi._Map)
    assert L.tbf_sizeof_midi_state() == C.sizeof(midi.MidiState)
    assert L.tbf_sizeof_programme_table() == C.sizeof(midi.ProgrammeTable)


def test_oracle_struct_sizes_match(ref):
    L = ref.lib()
    assert L.ref_sizeof_tonegen_tables() == C.sizeof(cabi.TonegenTables)
    assert L.ref_sizeof_whirl_tables() == C.sizeof(cabi.WhirlTables)
    assert L.ref_sizeof_state_export() == C.sizeof(cabi.StateExport)


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(engine.TbfError):
        engine.Context(0)


def test_product_never_imports_oracle():
    """The product package must not reference oracle/ (no CPU fallback)."""
    pkg = os.path.join(ROOT, "tunebfree_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "ref_engine" not in txt and "libtbf_ref" not in txt, f
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f


def _sass():
    import shutil
    import subprocess
    from tunebfree_b200 import engine
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    assert os.path.exists(engine.LIB_PATH), "libtbf_gpu.so is not built (python -c 'import __graft_entry__ as g; g.build()')"
    assert os.path.exists(cuobjdump), "cuobjdump (CUDA toolkit) not found: the SASS guards cannot run"
    return subprocess.run([cuobjdump, "-sass", engine.LIB_PATH], capture_output=True, text=True,
                          check=True).stdout


def _check_sass():
    sass = _sass()
    # The packed sums are FFMA2 used as an ADDER (product * 1.0 + sum, the 1.0 a run-time value:
    # tbf_kernels.cu add2_via_fma), each fed by its own packed multiply.  Had ptxas contracted a
    # multiply into the add (one rounding instead of two) the FMUL2 would be gone: there must be
    # one FMUL2 per FFMA2, and no packed add at all (a packed add fed by a packed multiply is what
    # ptxas contracts even under --fmad=false)
    n_mul = len(re.findall(r"\bFMUL2\b", sass))
    n_fma = len(re.findall(r"\bFFMA2\b", sass))
    assert n_mul > 0 and n_fma > 0 and n_mul >= n_fma, (n_mul, n_fma)
    assert not re.search(r"\bFADD2\b", sass)
    # scalar FFMA / DFMA only inside the division sequences ptxas expands (their multiplicand or
    # addend is RZ or the constant 1): none with three live register operands in a float kernel
    for line in sass.splitlines():
        m = re.search(r"\bFFMA\s+(\S+), (\S+), (\S+), (\S+) ;", line)
        if m:
            assert "RZ" in (m.group(2), m.group(3)), "a fused float multiply-add: " + line.strip()
    # stage T's bulk-copy variants stage the waves in shared memory by TMA (UBLKCP) with mbarrier
    # completion (SYNCS); stages R and D move their organ x sample tiles by 2-D tensor-map TMA
    # (UTMALDG loads, UTMASTG stores); stage T's default loads are 128-bit
    assert "UBLKCP" in sass and "SYNCS" in sass
    assert "UTMALDG" in sass and "UTMASTG" in sass
    assert re.search(r"\bLDG\.E\.128", sass)


def test_no_contracted_packed_fma_in_kernels():
    """The tone generator's sum multiplies packed FP32 pairs (FMUL2) and adds them with FFMA2 used
    as an adder (x * 1.0 + sum: exact product, one rounding = add.rn).  ptxas contracts a
    mul.f32x2 feeding an add.f32x2 into one FFMA2 even with --fmad=false, which would change the
    reference's rounding; the kernels are written so that it cannot (tbf_kernels.cu,
    add2_via_fma) and this test pins it: one FMUL2 per FFMA2, no FADD2, no scalar FFMA of three
    live operands anywhere in the library's SASS, and the TMA
    bulk copies (UBLKCP) with their mbarriers (SYNCS) that stage the wave tables.  Fails -- does
    not skip -- when the library or cuobjdump is missing."""
    _check_sass()


@pytest.mark.gpu
def test_sass_guards_on_the_gpu_box():
    """the same guards in the -m gpu set: what ran on the GPU is what was checked"""
    _check_sass()


def _random_event_list(rng, n_events, n_inst, n_frag):
    slot = np.sort(rng.randint(0, n_inst, n_events))
    frag = rng.randint(0, n_frag, n_events)
    order = np.lexsort((frag, slot))
    ev = np.zeros((n_events, 5), dtype=np.int32)
    ev[:, 0], ev[:, 1] = slot[order], frag[order]
    ev[:, 2] = cabi.EV_DRAWBAR
    return ev


@pytest.mark.parametrize("n_events", [0, 1, 1000, (1 << 18) - 1, (1 << 18) + 12345, 1500000])
def test_event_index_matches_numpy(n_events):
    """tbf_gpu_event_index (what tbf_gpu_batch_prepare runs on the host; from 2^18 events on in ranges
    on several threads): start[s] = first event of a slot >= s, for lists with empty slots at both ends."""
    rng = np.random.RandomState(n_events % 1000 + 1)
    n_inst, n_frag = 4099, 3750
    ev = _random_event_list(rng, n_events, n_inst, n_frag)
    if n_events > 10:
        ev = ev[(ev[:, 0] > 2) & (ev[:, 0] < n_inst - 3) & (ev[:, 0] != 2000)]   # slots nobody names
    start = engine.event_index(ev, n_inst, n_frag)
    want = np.searchsorted(ev[:, 0], np.arange(n_inst + 1), side="left").astype(np.uint64)
    assert np.array_equal(start, want)


def test_event_index_names_the_broken_rule():
    rng = np.random.RandomState(3)
    n_inst, n_frag = 64, 500
    for n_events in (200, 400000):          # one thread | several
        ev = _random_event_list(rng, n_events, n_inst, n_frag)
        engine.event_index(ev, n_inst, n_frag)
        for where in (0, n_events // 3, n_events - 1):
            bad = ev.copy()
            bad[where, 1] = n_frag                       # a fragment past the batch
            bad = bad[np.lexsort((bad[:, 1], bad[:, 0]))]
            with pytest.raises(engine.TbfError, match="out of range"):
                engine.event_index(bad, n_inst, n_frag)
        bad = ev.copy()
        bad[-1, 0] = n_inst                              # a slot past the batch
        with pytest.raises(engine.TbfError, match="out of range"):
            engine.event_index(bad, n_inst, n_frag)
        for where in (1, n_events // 2, n_events - 1):
            bad = ev.copy()
            k = where
            while k > 0 and (bad[k, 0], bad[k, 1]) == (bad[k - 1, 0], bad[k - 1, 1]):
                k -= 1
            if k == 0:
                continue
            bad[[k - 1, k]] = bad[[k, k - 1]]            # two neighbours swapped
            with pytest.raises(engine.TbfError, match="not sorted"):
                engine.event_index(bad, n_inst, n_frag)
