#!/usr/bin/env python
"""Generate tests/golden/*.npz from the ORACLE (the reference engine compiled in place,
oracle/_ref).  Run in the build container (needs /root/reference or a prebuilt oracle/_ref):

    python tests/golden/make_golden.py

Each file holds the rendered L/R float32 arrays (bit patterns are what the tests compare) of one
case of cases.py, plus a few integer state fields after the render."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_engine  # noqa: E402
from tunebfree_b200 import tables  # noqa: E402
from tests.golden.cases import CASES  # noqa: E402


def main():
    for name, c in CASES.items():
        freqs = tables.freqs_from_scl(c["scl"]) if c["scl"] else None
        r = ref_engine.RefInstance(freqs=freqs, seed=c["seed"], cfg=c["cfg"])
        l, rr = r.render(c["n_frag"], c["flags"], c["events"])
        st = r.export_state()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), l=l, r=rr,
                            active=np.int32(st.active_osc_l_end),
                            osc_pos=np.array(list(st.osc_pos), dtype=np.uint32),
                            whirl_outpos=np.uint32(st.whirl_outpos),
                            vib_out_pos=np.uint32(st.vib_out_pos))
        print("%-28s %6d samples, peak %.4f, %d active wheels" % (
            name, l.size, float(np.abs(l).max()), st.active_osc_l_end))
        r.close()


if __name__ == "__main__":
    main()
