/* test_multi.c -- a plain C host driving two devices through tbf_multi (include/tbf_multi.h):
 * 24 organs of three table sets (engine seeds 1, 2, 3), planned over two entries by table key,
 * rendered in one tbf_multi_render_batch, compared bit for bit with the same organs rendered on
 * ONE context.  Usage: test_multi <device A> <device B>   (A == B: two contexts on one GPU)
 * Exit 0 = identical; anything else prints why. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "tbf_gpu.h"
#include "tbf_multi.h"
#include "tbf_tables.h"

#define N_ORGANS 24
#define N_SETS 3
#define N_FRAG 48
#define FRAMES (N_FRAG * TBF_FRAGMENT)

#define CHECK(call)                                                      \
    do {                                                                 \
        int rc_ = (call);                                                \
        if (rc_ != TBF_OK) {                                             \
            fprintf(stderr, "%s -> %d (line %d)\n", #call, rc_, __LINE__); \
            return 2;                                                    \
        }                                                                \
    } while (0)

static int add_ev(tbf_event *ev, int n, unsigned slot, unsigned frag, int type, int a, int b)
{
    ev[n].instance = slot;
    ev[n].fragment = frag;
    ev[n].type = type;
    ev[n].a = a;
    ev[n].b = b;
    return n + 1;
}

/* a small chord script per slot: drawbars, vibrato C3, Leslie fast, keys on / off */
static int script(tbf_event *ev, int n, unsigned slot, unsigned organ)
{
    for (int d = 0; d < 9; ++d)
        n = add_ev(ev, n, slot, 0, TBF_EV_DRAWBAR, d, 8);
    n = add_ev(ev, n, slot, 0, TBF_EV_PERC_ENABLE, 1, 0);
    n = add_ev(ev, n, slot, 0, TBF_EV_VIB_UPPER, 1, 0);
    n = add_ev(ev, n, slot, 0, TBF_EV_VIB_KNOB, 5, 0);
    n = add_ev(ev, n, slot, 0, TBF_EV_REV_OPTION, 8, 0);
    for (int k = 0; k < 4; ++k)
        n = add_ev(ev, n, slot, 1, TBF_EV_KEY_ON, 40 + (int)((organ * 7 + k * 5) % 50), 0);
    for (int k = 0; k < 4; ++k)
        n = add_ev(ev, n, slot, 30, TBF_EV_KEY_OFF, 40 + (int)((organ * 7 + k * 5) % 50), 0);
    return n;
}

int main(int argc, char **argv)
{
    int devices[2] = {argc > 1 ? atoi(argv[1]) : 0, argc > 2 ? atoi(argv[2]) : 0};
    tbf_tonegen_tables tt[N_SETS];
    tbf_whirl_tables wt;
    tbf_whirl_spec ws;
    tbf_whirl_spec_default(&ws, 48000.0);
    CHECK(tbf_build_whirl_tables(&ws, &wt));
    for (int s = 0; s < N_SETS; ++s) {
        tbf_tonegen_spec ts;
        tbf_tonegen_spec_default(&ts, 48000.0);
        tbf_freqs_12tet(ts.freqs);
        ts.seed = 1 + (unsigned)s;
        CHECK(tbf_build_tonegen_tables(&ts, &tt[s]));
    }
    /* organ i plays with table set i % 3: the plan must bring each set's organs together */
    uint64_t key[N_ORGANS];
    uint32_t device_of[N_ORGANS], order[N_ORGANS];
    for (int i = 0; i < N_ORGANS; ++i)
        key[i] = 1000u + (uint64_t)(i % N_SETS);
    CHECK(tbf_multi_plan(2, N_ORGANS, key, device_of, order));
    for (int k = 1; k < N_ORGANS; ++k)
        if (key[order[k]] < key[order[k - 1]] || device_of[order[k]] < device_of[order[k - 1]]) {
            fprintf(stderr, "plan is not sorted by key / not a block partition\n");
            return 3;
        }

    /* ---- reference: one context, organs in plan order ---- */
    static float ref_l[N_ORGANS][FRAMES], ref_r[N_ORGANS][FRAMES], got_l[N_ORGANS][FRAMES], got_r[N_ORGANS][FRAMES];
    static tbf_event ev[N_ORGANS * 32];
    uint64_t ref_cs[N_ORGANS], got_cs[N_ORGANS];
    int n_ev = 0;
    for (int k = 0; k < N_ORGANS; ++k)
        n_ev = script(ev, n_ev, (unsigned)k, order[k]);
    {
        tbf_ctx *ctx;
        CHECK(tbf_gpu_create(&ctx, devices[0]));
        int tg[N_SETS], wh;
        uint32_t ids[N_ORGANS];
        CHECK(tbf_gpu_whirl_tables_create(ctx, &wt, &wh));
        for (int s = 0; s < N_SETS; ++s)
            CHECK(tbf_gpu_tonegen_tables_create(ctx, &tt[s], &tg[s]));
        for (int k = 0; k < N_ORGANS; ++k)
            CHECK(tbf_gpu_instance_create(ctx, tg[order[k] % N_SETS], wh, 1, &ids[k]));
        tbf_batch_desc d;
        memset(&d, 0, sizeof(d));
        d.n_instances = N_ORGANS;
        d.instance_ids = ids;
        d.n_fragments = N_FRAG;
        d.events = ev;
        d.n_events = (uint64_t)n_ev;
        d.out_l = &ref_l[0][0];
        d.out_r = &ref_r[0][0];
        d.out_stride = FRAMES;
        d.flags = TBF_RENDER_WHIRLPROC3;
        d.checksums = ref_cs;
        CHECK(tbf_gpu_render_batch(ctx, &d));
        tbf_gpu_destroy(ctx);
    }

    /* ---- the same through tbf_multi: every entry uploads only the sets its organs use ---- */
    tbf_multi *m;
    CHECK(tbf_multi_create(&m, 2, devices));
    if (tbf_multi_device_count(m) != 2)
        return 4;
    uint32_t inst[N_ORGANS], slot_dev[N_ORGANS];
    int uploads = 0;
    for (int e = 0; e < 2; ++e) {
        tbf_ctx *ctx = tbf_multi_context(m, e);
        int tg[N_SETS] = {-1, -1, -1}, wh = -1;
        for (int k = 0; k < N_ORGANS; ++k) {
            const int s = (int)(order[k] % N_SETS);
            if (device_of[order[k]] != (uint32_t)e)
                continue;
            if (wh < 0)
                CHECK(tbf_gpu_whirl_tables_create(ctx, &wt, &wh));
            if (tg[s] < 0) {
                CHECK(tbf_gpu_tonegen_tables_create(ctx, &tt[s], &tg[s]));
                ++uploads;
            }
            CHECK(tbf_gpu_instance_create(ctx, tg[s], wh, 1, &inst[k]));
            slot_dev[k] = (uint32_t)e;
        }
    }
    if (uploads > N_SETS + 1) { /* 3 sets over 2 devices: at most one set straddles the cut */
        fprintf(stderr, "plan scattered the table sets: %d uploads\n", uploads);
        return 5;
    }
    tbf_multi_batch b;
    memset(&b, 0, sizeof(b));
    b.n_slots = N_ORGANS;
    b.device_of = slot_dev;
    b.instance_ids = inst;
    b.n_fragments = N_FRAG;
    b.events = ev;
    b.n_events = (uint64_t)n_ev;
    b.out_l = &got_l[0][0];
    b.out_r = &got_r[0][0];
    b.out_stride = FRAMES;
    b.flags = TBF_RENDER_WHIRLPROC3;
    b.checksums = got_cs;
    float ms[2] = {0.f, 0.f};
    int rc = tbf_multi_render_batch(m, &b, ms);
    if (rc != TBF_OK) {
        fprintf(stderr, "tbf_multi_render_batch -> %d: %s\n", rc, tbf_multi_last_error(m));
        return 6;
    }
    tbf_multi_destroy(m);
    double peak = 0.0;
    for (int k = 0; k < N_ORGANS; ++k) {
        if (memcmp(got_l[k], ref_l[k], sizeof(ref_l[k])) || memcmp(got_r[k], ref_r[k], sizeof(ref_r[k])) ||
            got_cs[k] != ref_cs[k]) {
            fprintf(stderr, "slot %d (organ %u, entry %u) differs\n", k, order[k], slot_dev[k]);
            return 7;
        }
        for (int i = 0; i < FRAMES; ++i)
            if (ref_l[k][i] > peak)
                peak = ref_l[k][i];
    }
    if (peak < 0.01 || ms[0] <= 0.f || ms[1] <= 0.f) {
        fprintf(stderr, "silent render or missing device times (peak %g, %g ms, %g ms)\n", peak, ms[0], ms[1]);
        return 8;
    }
    for (int s = 0; s < N_SETS; ++s)
        tbf_free_tonegen_tables(&tt[s]);
    tbf_free_whirl_tables(&wt);
    printf("tbf_multi: %d organs over entries on devices %d and %d identical to one context (peak %.3f, %.2f / %.2f ms)\n",
           N_ORGANS, devices[0], devices[1], peak, ms[0], ms[1]);
    return 0;
}
