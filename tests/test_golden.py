"""Committed golden vectors (tests/golden/*.npz, generated from the reference engine by
tests/golden/make_golden.py).

* not gpu: the oracle reproduces them here (guards the oracle build / flags against drift);
* gpu: the CUDA path, fed by the product's OWN host-side table builder (no oracle, no
  /root/reference at run time), reproduces them bit for bit.
"""
import os

import numpy as np
import pytest

from tests.golden.cases import CASES
from tunebfree_b200 import tables

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_reproduces_golden(ref, name):
    c, g = CASES[name], _load(name)
    freqs = tables.freqs_from_scl(c["scl"]) if c["scl"] else None
    r = ref.RefInstance(freqs=freqs, seed=c["seed"], cfg=c["cfg"])
    l, rr = r.render(c["n_frag"], c["flags"], c["events"])
    st = r.export_state()
    r.close()
    assert np.array_equal(_bits(l), _bits(g["l"])) and np.array_equal(_bits(rr), _bits(g["r"]))
    assert st.active_osc_l_end == int(g["active"])
    assert np.array_equal(np.array(list(st.osc_pos), dtype=np.uint32), g["osc_pos"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_reproduces_golden(gpu_ctx, name):
    c, g = CASES[name], _load(name)
    freqs = tables.freqs_from_scl(c["scl"]) if c["scl"] else None
    spec = tables.InstanceSpec(freqs=freqs, seed=c["seed"], cfg=c["cfg"])
    tt, wt = spec.build()
    tg = gpu_ctx.tonegen_tables(tt)
    wh = gpu_ctx.whirl_tables(wt)
    tables.free_tables(tt, wt)
    first = gpu_ctx.create_instances(tg, wh, 3)
    # three copies of the organ in one batch: every slot must give the same audio
    l, r = gpu_ctx.render(3, c["n_frag"], [c["events"]] * 3, c["flags"],
                          ids=[first, first + 1, first + 2])
    for s in range(3):
        bad = np.nonzero(_bits(l[s]) != _bits(g["l"]))[0]
        assert bad.size == 0, "%s slot %d: L differs first at sample %d" % (name, s, bad[0])
        assert np.array_equal(_bits(r[s]), _bits(g["r"]))
    st = gpu_ctx.export_state(first + 2)
    assert st.active_osc_l_end == int(g["active"])
    assert np.array_equal(np.array(list(st.osc_pos), dtype=np.uint32), g["osc_pos"])
    assert st.whirl_outpos == int(g["whirl_outpos"]) and st.vib_out_pos == int(g["vib_out_pos"])
    assert st.error_flags == 0
