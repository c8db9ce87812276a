"""CPU tests of the host-side table builder (tbf_build_*_tables, the product's own
restatement of initToneGenerator / init_vibrato / initWhirl) against the oracle's
struct dumps: every table entry must be BIT-identical."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from tunebfree_b200 import cabi, tables

REF = "/root/reference"


def _struct_diffs(a, b):
    bad = []
    for fname, _ in a._fields_:
        va, vb = getattr(a, fname), getattr(b, fname)
        if isinstance(va, C._Pointer):
            continue
        if isinstance(va, C.Array):
            if bytes(va) != bytes(vb):
                bad.append(fname)
        elif isinstance(va, float):
            if np.float64(va).tobytes() != np.float64(vb).tobytes():
                bad.append((fname, va, vb))
        elif va != vb:
            bad.append((fname, va, vb))
    return bad


def _ptr_equal(a, b, n, dtype):
    xa = np.ctypeslib.as_array(a, shape=(n,)).view(dtype)
    xb = np.ctypeslib.as_array(b, shape=(n,)).view(dtype)
    return np.array_equal(xa, xb)


def assert_tables_identical(ref, seed=1, cfg="", freqs=None, target_ratio=None):
    r = ref.RefInstance(freqs=freqs, seed=seed, cfg=cfg, target_ratio=target_ratio)
    rt, rw = r.tonegen_tables(), r.whirl_tables()
    spec = tables.InstanceSpec(freqs=freqs if freqs is not None else ref.tet12(), seed=seed, cfg=cfg,
                               target_ratio=target_ratio)
    tt, wt = spec.build()
    try:
        assert _struct_diffs(tt, rt) == []
        assert tt.wave_total == rt.wave_total
        assert _ptr_equal(tt.wave_data, rt.wave_data, tt.wave_total, np.uint32)
        nc = rt.contrib_start[cabi.MAX_KEYS]
        assert tt.contrib_start[cabi.MAX_KEYS] == nc
        assert _ptr_equal(tt.contrib_wheel, rt.contrib_wheel, nc, np.uint16)
        assert _ptr_equal(tt.contrib_bus, rt.contrib_bus, nc, np.uint16)
        assert _ptr_equal(tt.contrib_level, rt.contrib_level, nc, np.uint32)
        assert _struct_diffs(wt, rw) == []
        N = cabi.WHIRL_DISPLC
        for name, k in (("hn_fwd_displ", 1), ("hn_bwd_displ", 1), ("dr_fwd_displ", 1),
                        ("dr_bwd_displ", 1), ("bfw", 5), ("bbw", 5)):
            assert _ptr_equal(getattr(wt, name), getattr(rw, name), N * k, np.uint32), name
    finally:
        tables.free_tables(tt, wt)
        r.close()
    return spec


def test_glibc_rand_stream():
    """tbf_rand reproduces glibc's rand() after srand(seed) (TYPE_3 generator)."""
    libc = C.CDLL("libc.so.6")
    st = (C.c_int32 * 35)()
    L = tables._lib()
    for seed in (0, 1, 2, 12345, 0x7FFFFFFF, 0xFFFFFFFF):
        libc.srand(C.c_uint(seed))
        L.tbf_srand(st, seed)
        for _ in range(2000):
            assert libc.rand() == L.tbf_rand(st)
    libc.srand(1)
    assert [libc.rand() for _ in range(3)] == [1804289383, 846930886, 1681692777]


def test_default_tables_identical(ref):
    spec = assert_tables_identical(ref, seed=1)
    assert spec.rejected == [] and spec.unclaimed == []


@pytest.mark.parametrize("seed", [2, 16, 99])
def test_seeds(ref, seed):
    assert_tables_identical(ref, seed=seed)


CFGS = [
    "osc.harmonic.1=1.0\nosc.harmonic.2=0.5\nosc.harmonic.3=0.333\nosc.harmonic.4=0.25\n"
    "osc.harmonic.5=0.2\nosc.harmonic.6=0.166\nosc.harmonic.13=0.1\n",
    "osc.compartment-crosstalk=0\nosc.transformer-crosstalk=0\nosc.terminalstrip-crosstalk=0\n"
    "osc.wiring-crosstalk=0\nosc.contribution-floor=0.001\nosc.contribution-min=0.002\n",
    "osc.compartment-crosstalk=0.05\nosc.terminalstrip-crosstalk=0.03\nosc.wiring-crosstalk=0.2\n",
    "osc.eq.macro=peak24\n",
    "osc.eq.macro=peak46\nosc.x-precision=0.01\n",
    "osc.eq.p1y=0.3\nosc.eq.r1y=1.2\nosc.eq.p4y=0.9\nosc.eq.r4y=-1.5\n",
    "osc.attack.model=cosine\nosc.release.model=click\nosc.release.click.level=0.7\n",
    "osc.attack.model=shelf\nosc.release.model=shelf\n",
    "osc.attack.model=linear\nosc.release.model=cosine\n",
    "osc.attack.click.level=0.9\nosc.attack.click.minlength=0.1\nosc.attack.click.maxlength=0.6\n",
    "osc.attack.click.minlength=0.5\nosc.attack.click.maxlength=0.2\n",
    "osc.perc.normal=0.8\nosc.perc.soft=0.3\nosc.perc.gain=7.5\nosc.perc.bus.a=2\nosc.perc.bus.b=5\n"
    "osc.perc.bus.trig=-1\nosc.perc.fast=0.5\nosc.perc.slow=3\n",
    "osc.perc.bus.trig=2\n",
    "scanner.hz=6.1\nscanner.modulation.v1=2.5\nscanner.modulation.v2=7\nscanner.modulation.v3=11\n",
    "scanner.hz=30\nosc.wiring-crosstalk=2\nwhirl.bypass=4\n",  # all out of range: ignored
    "whirl.horn.filter.b.gain=-60\nwhirl.horn.filter.a.hz=3000\nwhirl.horn.filter.a.q=1.2\n"
    "whirl.horn.filter.a.type=6\nwhirl.horn.filter.a.gain=4\n",
    "whirl.drum.filter.type=0\nwhirl.drum.filter.hz=700\nwhirl.drum.filter.q=0.9\nwhirl.drum.filter.gain=-3\n",
    "whirl.horn.brakepos=0.25\nwhirl.drum.brakepos=0.5\nwhirl.speed-preset=2\n",
    "whirl.horn.mic.angle=90\nwhirl.drum.width=0.4\nwhirl.horn.width=-0.3\nwhirl.bypass=1\n",
    "whirl.horn.slowrpm=30\nwhirl.horn.fastrpm=500\nwhirl.drum.slowrpm=20\nwhirl.drum.fastrpm=300\n"
    "whirl.horn.acceleration=0.3\nwhirl.horn.deceleration=0.5\nwhirl.drum.acceleration=2\n"
    "whirl.drum.deceleration=1\nwhirl.speed-preset=1\n",
    "whirl.horn.radius=25\nwhirl.drum.radius=30\nwhirl.mic.distance=60\nwhirl.horn.offset.x=3\n"
    "whirl.horn.offset.z=-2\nwhirl.horn.level=0.5\nwhirl.horn.leak=0.3\n",
    "whirl.horn.comb.a.feedback=-0.55\nwhirl.horn.comb.a.delay=38\n",
]


@pytest.mark.parametrize("i", range(len(CFGS)))
def test_cfg_variants(ref, i):
    assert_tables_identical(ref, seed=3 + i, cfg=CFGS[i])


def test_transformer_crosstalk_is_reference_abort():
    """osc.transformer-crosstalk > 0 makes the reference assert(0) for wheels > 91
    (src/tonegen.cpp:925-937): reported as TBF_ERR_REFERENCE_UB, not rendered."""
    spec = tables.InstanceSpec(cfg="osc.transformer-crosstalk=0.01\n")
    with pytest.raises(ValueError):
        spec.build()


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")
@pytest.mark.parametrize("name", ["osc-bassy", "osc-middy", "osc-noxtalk", "osc-sawtooth", "osc-scoop",
                                  "osc-square", "osc-trebly", "osc-triangle", "osc-xwave",
                                  "whirl-mellow", "whirl-sharp", "whirl-x100"])
def test_reference_cfg_files(ref, name):
    """BASELINE config 5's cfg/osc-*.cfg and cfg/whirl-*.cfg variants."""
    text = open(os.path.join(REF, "cfg", name + ".cfg"), encoding="latin-1").read()
    text = "\n".join(l for l in text.splitlines() if not l.strip().lower().startswith("config.read"))
    spec = assert_tables_identical(ref, seed=4, cfg=text)
    assert spec.rejected == []


def test_sample_rates(ref):
    for rate in (44100.0, 96000.0):
        r = ref.RefInstance(seed=1, rate=rate)
        spec = tables.InstanceSpec(freqs=ref.tet12(), seed=1, rate=rate)
        tt, wt = spec.build()
        rt, rw = r.tonegen_tables(), r.whirl_tables()
        assert _struct_diffs(tt, rt) == [] and _struct_diffs(wt, rw) == []
        assert tt.wave_total == rt.wave_total
        assert _ptr_equal(tt.wave_data, rt.wave_data, tt.wave_total, np.uint32)
        tables.free_tables(tt, wt)
        r.close()


def test_target_ratios(ref):
    assert_tables_identical(ref, seed=2, target_ratio=[0.5, 1.5, 1, 2, 3, 4, 5, 7, 8])


SCL_19 = "! 19tet\n19-TET\n 19\n!\n" + "\n".join("%.5f" % (1200.0 * i / 19) for i in range(1, 19)) + "\n2/1\n"
SCL_BP = "! bp\nBohlen-Pierce\n 13\n!\n" + "\n".join("%.5f" % (1901.955 * i / 13) for i in range(1, 13)) + "\n3/1\n"
SCL_JI = "! ji\n5-limit\n 7\n!\n9/8\n5/4\n4/3\n3/2\n5/3\n15/8\n2\n"


@pytest.mark.parametrize("scl", [SCL_19, SCL_BP, SCL_JI])
def test_microtunings(ref, scl):
    """Per-instance tuning -> wheel choice, loop lengths and wavetables (BASELINE config 4)."""
    freqs = tables.freqs_from_scl(scl)
    assert_tables_identical(ref, seed=1, freqs=freqs)


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "tests/regression_test_data")),
                    reason="reference tree not present")
def test_regression_fixture_tunings(ref):
    """The .scl files of the reference's own regression fixtures."""
    for path in sorted(glob.glob(os.path.join(REF, "tests/regression_test_data/*/*.scl"))):
        freqs = tables.freqs_from_scl(open(path, encoding="latin-1").read())
        assert_tables_identical(ref, seed=1, freqs=freqs)


def test_extend_frequencies_known_answers():
    """Known answers of the reference's own doctest cases (src/tuning.cpp:189-207):
    12-TET extends by octaves: f[128] = 2 * f[116]."""
    ext, ss, per = tables.extend_frequencies(tables.freqs_12tet())
    assert (ss, per) == (12, 2.0)
    assert ext[128] == 2.0 * ext[116] and ext[299] == 2.0 * ext[287]
    f19 = tables.freqs_from_scl(SCL_19)
    ext, ss, per = tables.extend_frequencies(f19)
    assert (ss, per) == (19, 2.0)
    ext, ss, per = tables.extend_frequencies(tables.freqs_from_scl(SCL_BP))
    assert (ss, per) == (13, 3.0)
