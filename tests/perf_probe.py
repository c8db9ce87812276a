"""Developer probe (not a bench line): device time of the render kernel with tables
taken from the oracle.  Usage: python tests/perf_probe.py [n_instances] [n_fragments] [config]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from oracle import ref_engine
from tunebfree_b200 import cabi, engine, workloads

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 375
cfg = int(sys.argv[3]) if len(sys.argv) > 3 else 3
r = ref_engine.RefInstance(seed=1)
ctx = engine.Context(0)
tg = ctx.tonegen_tables(r.tonegen_tables())
wh = ctx.whirl_tables(r.whirl_tables())
first = ctx.create_instances(tg, wh, n)
ev = engine.events_from_numpy(workloads.batch_events_numpy(0, n, nf, cfg))
flags = workloads.render_flags(cfg)
for it in range(3):
    cs = None if os.environ.get("PROBE_NO_CSUM") else np.zeros(n, dtype=np.uint64)
    t0 = time.time()
    ctx.render_device(n, nf, ev, flags, None, None, nf * 128, checksums=cs)
    wall = time.time() - t0
    ms, _ = ctx.last_launch_ms()
    secs = n * nf / 375.0
    print("iter %d: kernel %.2f ms (wall %.2f ms) -> %.0f organ-s/s  [%d inst x %d frag, config %d]" % (
        it, ms, wall * 1e3, secs / (ms * 1e-3), n, nf, cfg))
if os.environ.get("PROBE_STAGES"):
    ctx.set_stage_profiling(True)
    ctx.render_device(n, nf, ev, flags, None, None, nf * 128)
    st = ctx.stage_times()
    ctx.set_stage_profiling(False)
    print("stages: " + "  ".join("%s %.2f ms/%d" % (k, v[0], v[1]) for k, v in st.items()))
