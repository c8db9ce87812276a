"""BASELINE config 4: one organ per Scala scale of the reference's archive
(/root/reference/tests/scala_scale_archive/scl, 5176 files).

tunebfree_b200/data/scala_sweep.npz (tests/golden/make_scala_sweep.py, run in the build container
against the reference engine) holds, for every file, the scale's degrees, what each side did
with it and a digest of the REFERENCE's tone-generator tables.  The CPU tests rebuild a sample
of the tunings with the product's table builder and compare digests; the GPU test renders a
sample through config 3's chain and compares with the oracle bit for bit."""
import ctypes as C
import os

import numpy as np
import pytest

from tunebfree_b200 import cabi, tables, workloads

FIXTURE = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tunebfree_b200", "data",
                       "scala_sweep.npz")
ARCHIVE = "/root/reference/tests/scala_scale_archive/scl"


@pytest.fixture(scope="module")
def sweep():
    return np.load(FIXTURE)


def _degrees(z, i):
    return z["degrees"][z["deg_off"][i]:z["deg_off"][i + 1]]


def test_fixture_census(sweep):
    st = sweep["status"]
    assert len(st) == 5176
    assert int((st == 3).sum()) == 0, "table mismatch against the reference"
    assert int((st == 4).sum()) == 0, "builder refuses a tuning the reference accepts"
    assert int((st == 0).sum()) > 4500
    # every tuning the reference aborts on is refused by the builder with an error code
    for i in np.flatnonzero(st == 2):
        assert sweep["ours_rc"][i] in (cabi.ERR_REFERENCE_UB, cabi.ERR_ARG), sweep["names"][i]


def test_builder_matches_reference_digest(sweep):
    """Every 37th accepted tuning: tables built here hash to what the reference's hashed to."""
    ok = np.flatnonzero(sweep["status"] == 0)[::37]
    assert len(ok) > 100
    for i in ok:
        spec = tables.InstanceSpec(freqs=tables.freqs_from_ratios(_degrees(sweep, i)), seed=1)
        tt = spec.build_tonegen()
        try:
            got = tables.table_digest(tt)
        finally:
            tables.free_tables(tt, None)
        assert got == tuple(int(x) for x in sweep["digest"][i]), sweep["names"][i]


def test_refused_tunings_do_not_build(sweep):
    for i in np.flatnonzero(sweep["status"] == 2)[:20]:
        spec = tables.InstanceSpec(freqs=tables.freqs_from_ratios(_degrees(sweep, i)), seed=1)
        tt = cabi.TonegenTables()
        rc = tables._lib().tbf_build_tonegen_tables(C.byref(spec.tg), C.byref(tt))
        assert rc == sweep["ours_rc"][i] and rc != cabi.OK


@pytest.mark.skipif(not os.path.isdir(ARCHIVE), reason="reference tree absent")
def test_parser_reproduces_fixture_degrees(sweep):
    names = sweep["names"]
    for i in range(0, len(names), 11):
        text = open(os.path.join(ARCHIVE, str(names[i])), encoding="latin-1").read()
        if sweep["status"][i] == 1:
            with pytest.raises(ValueError):
                tables.scl_ratios(text)
            continue
        deg = tables.scl_ratios(text)
        assert deg.tobytes() == _degrees(sweep, i).tobytes(), names[i]
        assert np.array_equal(tables.freqs_from_scl(text), tables.freqs_from_ratios(deg))


def test_workload_lists_accepted_scales(sweep):
    names, degs, digest = workloads.scala_sweep()
    assert len(names) == int((sweep["status"] == 0).sum()) == len(degs) == len(digest)
    ev = workloads.instance_events(7, 750, config=4)
    keys = {e[2] for e in ev if e[1] == cabi.EV_KEY_ON}
    assert keys == set(workloads.SWEEP_CHORD)


@pytest.mark.gpu
def test_gpu_sweep_sample_bit_exact():
    """48 tunings spread over the archive (5- to 100+-tone, octave and non-octave), each its own
    table set, 1 s through tonegen + vibrato C3 + Leslie fast: bit-identical to the oracle."""
    from oracle import ref_engine
    from tunebfree_b200 import engine
    names, degs, _ = workloads.scala_sweep()
    pick = list(range(0, len(names), len(names) // 47))[:48]
    n_frag = 375
    ev = workloads.instance_events(0, n_frag, config=4)
    ctx = engine.Context(0)
    tt0, wt0 = tables.InstanceSpec(seed=1).build()
    wh = ctx.whirl_tables(wt0)
    tables.free_tables(tt0, wt0)
    ids = []
    for g in pick:
        tt = tables.InstanceSpec(freqs=tables.freqs_from_ratios(degs[g]), seed=1).build_tonegen()
        ids.append(ctx.create_instances(ctx.tonegen_tables(tt), wh, 1))
        tables.free_tables(tt, None)
    gl, gr = ctx.render(len(pick), n_frag, [ev] * len(pick), cabi.RENDER_WHIRLPROC3, ids=ids)
    ctx.close()
    peaks = []
    for k, g in enumerate(pick):
        r = ref_engine.RefInstance(freqs=tables.freqs_from_ratios(degs[g]), seed=1)
        rl, rr = r.render(n_frag, cabi.RENDER_WHIRLPROC3, ev)
        r.close()
        assert np.array_equal(gl[k].view(np.uint32), rl.view(np.uint32)), names[g]
        assert np.array_equal(gr[k].view(np.uint32), rr.view(np.uint32)), names[g]
        peaks.append(float(np.abs(rl).max()))
    assert min(peaks) > 0.01
