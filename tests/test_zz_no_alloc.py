"""tbf_gpu_render_small_noalloc (include/tbf_gpu.h): the render chain with every intermediate
buffer in the library's static device array -- no counterpart in the reference, added
for callers that may not call cudaMalloc.

Two calls in a row (the second continues the organs' state and reuses the array with the first
call's contents still in it) must be bit-identical -- the tolerance of every parity test here --
to tbf_gpu_render_batch on twin organs fed the same events; batches beyond the static limits
are rejected with a message naming them.  Collected last on purpose
(file name)."""
import numpy as np
import pytest

from tunebfree_b200 import cabi, engine, tables, workloads

pytestmark = pytest.mark.gpu


def _split(ev, lo, hi):
    return [(f - lo, t, a, b) for (f, t, a, b) in ev if lo <= f < hi]


@pytest.mark.parametrize("flags", [cabi.RENDER_WHIRLPROC3, cabi.RENDER_WHIRL_BYPASS, 0])
def test_no_alloc_entry_matches_render_batch_on_two_calls_in_a_row(flags):
    import torch
    n, nf = 7, 32
    ctx = engine.Context(0)
    tt, wt = tables.InstanceSpec(seed=3).build()
    tg, wh = ctx.tonegen_tables(tt), ctx.whirl_tables(wt)
    tables.free_tables(tt, wt)
    a = ctx.create_instances(tg, wh, n)   # rendered by the existing entry point
    b = ctx.create_instances(tg, wh, n)   # rendered by the no-alloc entry point
    evs = [workloads.instance_events(10 + g, 2 * nf, config=3, restrike=20, hold=12) for g in range(n)]
    frames = nf * 128
    for call in range(2):
        part = [_split(e, call * nf, (call + 1) * nf) for e in evs]
        want = torch.zeros((2, n, frames), dtype=torch.float32, device="cuda:0")
        got = torch.full((2, n, frames), 7.0, dtype=torch.float32, device="cuda:0")
        cs_want = np.zeros(n, dtype=np.uint64)
        cs_got = np.zeros(n, dtype=np.uint64)
        ctx.render_device(n, nf, part, flags, want[0].data_ptr(), want[1].data_ptr(), frames,
                          ids=list(range(a, a + n)), checksums=cs_want)
        torch.cuda.synchronize()
        before = torch.cuda.mem_get_info(0)[0]
        ctx.render_small_noalloc(n, nf, part, flags, got[0].data_ptr(), got[1].data_ptr(), frames,
                                 ids=list(range(b, b + n)), checksums=cs_got)
        torch.cuda.synchronize()
        # the process's first call on a device makes the CUDA runtime load the module that holds
        # the static array (lazy module loading); from then on device memory does not move
        if call == 1:
            assert torch.cuda.mem_get_info(0)[0] == before, "the no-alloc entry point changed device memory"
        w, g = want.cpu().numpy(), got.cpu().numpy()
        assert float(np.abs(w).max()) > 0.01
        assert np.array_equal(w.view(np.uint32), g.view(np.uint32)), (flags, call)
        assert np.array_equal(cs_want, cs_got), (flags, call)
    # the organs' state is where the existing entry point leaves its twins
    sa, sb = ctx.export_state(a + 2), ctx.export_state(b + 2)
    assert list(sa.osc_pos) == list(sb.osc_pos) and sa.whirl_outpos == sb.whirl_outpos
    assert sa.vib_stator == sb.vib_stator and sa.horn_angle == sb.horn_angle
    ctx.close()


def test_no_alloc_entry_rejects_what_does_not_fit():
    import torch
    ctx = engine.Context(0)
    tt, wt = tables.InstanceSpec(seed=1).build()
    tg, wh = ctx.tonegen_tables(tt), ctx.whirl_tables(wt)
    tables.free_tables(tt, wt)
    first = ctx.create_instances(tg, wh, 40)
    buf = torch.zeros((2, 40, 64 * 128), dtype=torch.float32, device="cuda:0")
    lib = engine.lib()
    assert 4 << 20 < lib.tbf_gpu_noalloc_scratch_bytes() < 16 << 20
    for n, nf in ((33, 8), (4, 33)):
        with pytest.raises(engine.TbfError) as e:
            ctx.render_small_noalloc(n, nf, [[]] * n, 0, buf[0].data_ptr(), buf[1].data_ptr(), 64 * 128,
                                     ids=list(range(first, first + n)))
        assert "max 32 instances x 32 fragments" in str(e.value)
    with pytest.raises(engine.TbfError):   # host output is the allocating entry point's business
        d, keep = ctx._desc(2, 4, [[]] * 2, 0, [first, first + 1], buf[0].data_ptr(), buf[1].data_ptr(), 512, None)
        ctx._ck(lib.tbf_gpu_render_small_noalloc(ctx._h, __import__("ctypes").byref(d)))
    # a fitting call still works afterwards
    ctx.render_small_noalloc(2, 4, [[(0, cabi.EV_KEY_ON, 60, 0)]] * 2, 0, buf[0].data_ptr(), buf[1].data_ptr(),
                             64 * 128, ids=[first, first + 1])
    assert float(buf[0, 0, :512].abs().max().item()) > 0.0
    ctx.close()
