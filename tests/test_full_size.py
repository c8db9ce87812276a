"""Full-size runs (BASELINE.json configs 2 and 3 instance counts) checked through properties
that do not need the audio on the host: a position-keyed checksum of every organ's L/R bit
patterns is formed on the device; sampled organs are compared with the oracle's render of the
same events through the same checksum ("a checksum of checksums"), the run is repeated after a
reset (determinism over the whole batch), and no organ may raise an error flag."""
import numpy as np
import pytest

from tunebfree_b200 import cabi, engine, workloads

pytestmark = pytest.mark.gpu


def _run(ref, ctx, n, n_frag, config, sampled):
    r0 = ref.RefInstance(seed=1)
    tg = ctx.tonegen_tables(r0.tonegen_tables())
    wh = ctx.whirl_tables(r0.whirl_tables())
    r0.close()
    first = ctx.create_instances(tg, wh, n)
    ids = list(range(first, first + n))
    ev_np = workloads.batch_events_numpy(0, n, n_frag, config)
    ev = engine.events_from_numpy(ev_np)
    flags = workloads.render_flags(config)
    cs = np.zeros(n, dtype=np.uint64)
    ctx.render_device(n, n_frag, ev, flags, None, None, n_frag * 128, ids=ids, checksums=cs)
    assert len(set(cs.tolist())) > n * 0.99          # different chords -> different audio
    for i in sampled:
        r = ref.RefInstance(seed=1)
        l, rr = r.render(n_frag, flags, workloads.instance_events(i, n_frag, config))
        r.close()
        assert np.abs(l).max() > 0.01
        assert engine.checksum_host(l, rr) == int(cs[i]), i
    for i in (0, n // 2, n - 1):
        assert ctx.export_state(first + i).error_flags == 0
    ctx.reset(first, n)
    cs2 = np.zeros(n, dtype=np.uint64)
    ctx.render_device(n, n_frag, ev, flags, None, None, n_frag * 128, ids=ids, checksums=cs2)
    assert np.array_equal(cs, cs2)


def test_config3_4096_organs(ref, gpu_ctx):
    """4096 organs x 1 s through tonegen + vibrato + Leslie (whirlProc3)."""
    _run(ref, gpu_ctx, 4096, 375, 3, sampled=(0, 1, 31, 32, 1000, 2047, 2048, 4095))


def test_config2_1024_organs(ref, gpu_ctx):
    """1024 organs x 2 s, tone generator only."""
    _run(ref, gpu_ctx, 1024, 750, 2, sampled=(0, 5, 511, 1023))
