// tbf_shim.cpp -- the reference's per-instance engine API on top of the batched CUDA renderer
// (include/tbf_shim.h).  Host code only: every call ends in tbf_gpu_render_batch on a
// process-wide context; there is no CPU rendering here and no fallback -- when the context
// cannot be created the audio calls write nothing and tbf_shim_status() says why.
#include "tbf_shim.h"

#include <cstdio>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "tbf_gpu.h"
#include "tbf_tables.h"

namespace {

struct Shim {
    tbf_ctx *ctx = nullptr;
    int status = 0;
    std::string error;
    int default_tg = -1, default_wh = -1; // the half a one-sided organ does not use
    int live = 0;
};
Shim g;

void fail(int rc, const char *what)
{
    if (g.status == 0) {
        g.status = rc;
        g.error = std::string(what) + ": " +
                  (g.ctx && rc == TBF_ERR_CUDA ? tbf_gpu_last_error(g.ctx) : "tbf_status " + std::to_string(rc));
        fprintf(stderr, "tbf_shim: %s\n", g.error.c_str());
    }
}

tbf_ctx *context()
{
    if (!g.ctx && g.status == 0) {
        const char *dev = getenv("TBF_SHIM_DEVICE");
        const int rc = tbf_gpu_create(&g.ctx, dev ? atoi(dev) : 0);
        if (rc != TBF_OK) {
            g.ctx = nullptr;
            fail(rc, "tbf_gpu_create");
        }
    }
    return g.ctx;
}

typedef std::vector<std::pair<std::string, std::string>> CfgLines;

// -1 bad value, 0 not of this family, 1 stored (the reference's *Config return convention)
int store_cfg(CfgLines &lines, const char *family, ConfigContext *cfg)
{
    if (!cfg || !cfg->name || !cfg->value || strncmp(cfg->name, family, strlen(family)) != 0)
        return 0;
    tbf_tonegen_spec ts;
    tbf_whirl_spec ws;
    tbf_tonegen_spec_default(&ts, 48000.0);
    tbf_whirl_spec_default(&ws, 48000.0);
    const int rc = tbf_spec_apply_cfg(&ts, &ws, cfg->name, cfg->value);
    if (rc == 1)
        lines.emplace_back(cfg->name, cfg->value);
    return rc < 0 ? -1 : rc;
}

int default_whirl_tables(tbf_ctx *ctx)
{
    if (g.default_wh < 0) {
        tbf_whirl_spec ws;
        tbf_whirl_spec_default(&ws, 48000.0);
        tbf_whirl_tables wt;
        int rc = tbf_build_whirl_tables(&ws, &wt);
        if (rc == TBF_OK) {
            rc = tbf_gpu_whirl_tables_create(ctx, &wt, &g.default_wh);
            tbf_free_whirl_tables(&wt);
        }
        if (rc != TBF_OK)
            fail(rc, "default Leslie tables");
    }
    return g.default_wh;
}

int default_tonegen_tables(tbf_ctx *ctx)
{
    if (g.default_tg < 0) {
        tbf_tonegen_spec ts;
        tbf_tonegen_spec_default(&ts, 48000.0);
        tbf_freqs_12tet(ts.freqs);
        ts.seed = 1;
        tbf_tonegen_tables tt;
        int rc = tbf_build_tonegen_tables(&ts, &tt);
        if (rc == TBF_OK) {
            rc = tbf_gpu_tonegen_tables_create(ctx, &tt, &g.default_tg);
            tbf_free_tonegen_tables(&tt);
        }
        if (rc != TBF_OK)
            fail(rc, "default tone generator tables");
    }
    return g.default_tg;
}

struct Organ {
    bool ready = false;
    uint32_t id = 0;
    int own_tables = -1; // the table set this half uploaded (the other half is a shared default)
    std::vector<tbf_event> pending; // setter calls since the last rendered fragment, in order
    std::vector<float> scratch;
    void push(int type, int a, int b = 0)
    {
        tbf_event e;
        e.instance = 0;
        e.fragment = 0;
        e.type = type;
        e.a = a;
        e.b = b;
        pending.push_back(e);
    }
    // one render call: `n_frag` fragments, the pending events before the first.  A host never
    // polls tbf_shim_status(): on any failure the caller's buffers are silenced, not left as
    // they were
    void render(uint32_t n_frag, uint32_t flags, const float *in, float *out_l, float *out_r,
                float *const *stems = nullptr)
    {
        tbf_ctx *ctx = context();
        auto silence = [&]() {
            const size_t bytes = sizeof(float) * (size_t)n_frag * TBF_FRAGMENT;
            if (out_l)
                memset(out_l, 0, bytes);
            if (out_r)
                memset(out_r, 0, bytes);
            for (int j = 0; stems && j < 4; ++j)
                if (stems[j])
                    memset(stems[j], 0, bytes);
        };
        if (!ctx || !ready) {
            silence();
            pending.clear();
            return;
        }
        tbf_batch_desc d;
        memset(&d, 0, sizeof(d));
        d.n_instances = 1;
        d.instance_ids = &id;
        d.n_fragments = n_frag;
        d.events = pending.data();
        d.n_events = pending.size();
        d.out_l = out_l;
        d.out_r = out_r;
        d.out_stride = (uint64_t)n_frag * TBF_FRAGMENT;
        d.flags = flags;
        d.in = in;
        d.in_stride = in ? d.out_stride : 0;
        if (stems) {
            d.flags |= TBF_RENDER_STEMS;
            for (int j = 0; j < 4; ++j)
                d.stems[j] = stems[j];
        }
        const int rc = tbf_gpu_render_batch(ctx, &d);
        if (rc != TBF_OK) {
            fail(rc, "tbf_gpu_render_batch");
            silence();
        }
        pending.clear();
    }
};

} // namespace

struct b_tonegen {
    Organ organ;
    CfgLines cfg;
    double freqs[128];
    unsigned seed = 1;
};

struct b_whirl {
    Organ organ;
    CfgLines cfg;
};

extern "C" {

/* ---------------- tone generator ---------------- */

struct b_tonegen *allocTonegen(void)
{
    b_tonegen *t = new b_tonegen();
    tbf_freqs_12tet(t->freqs);
    ++g.live;
    return t;
}

void tbf_shim_set_tuning(struct b_tonegen *t, const double freqs[128])
{
    memcpy(t->freqs, freqs, sizeof(t->freqs));
}

void tbf_shim_set_seed(struct b_tonegen *t, unsigned seed) { t->seed = seed; }

int oscConfig(struct b_tonegen *t, ConfigContext *cfg) { return store_cfg(t->cfg, "osc.", cfg); }

int scannerConfig(void *t, ConfigContext *cfg)
{
    return store_cfg(((b_tonegen *)t)->cfg, "scanner.", cfg);
}

void initToneGenerator(struct b_tonegen *t, void *, double rate, double *targetRatio)
{
    tbf_ctx *ctx = context();
    if (!ctx)
        return;
    tbf_tonegen_spec ts;
    tbf_whirl_spec ws;
    tbf_tonegen_spec_default(&ts, rate);
    tbf_whirl_spec_default(&ws, rate);
    memcpy(ts.freqs, t->freqs, sizeof(ts.freqs));
    if (targetRatio)
        memcpy(ts.target_ratio, targetRatio, sizeof(ts.target_ratio));
    ts.seed = t->seed;
    for (const auto &kv : t->cfg)
        tbf_spec_apply_cfg(&ts, &ws, kv.first.c_str(), kv.second.c_str());
    tbf_tonegen_tables tt;
    int rc = tbf_build_tonegen_tables(&ts, &tt);
    if (rc != TBF_OK) {
        fail(rc, "tbf_build_tonegen_tables"); /* the reference asserts here */
        return;
    }
    int tg = -1;
    rc = tbf_gpu_tonegen_tables_create(ctx, &tt, &tg);
    tbf_free_tonegen_tables(&tt);
    const int wh = rc == TBF_OK ? default_whirl_tables(ctx) : -1;
    if (rc == TBF_OK && wh >= 0)
        rc = tbf_gpu_instance_create(ctx, tg, wh, 1, &t->organ.id);
    if (rc != TBF_OK || wh < 0) {
        fail(rc, "tone generator upload");
        if (tg >= 0)
            tbf_gpu_tonegen_tables_destroy(ctx, tg); /* uploaded, but no organ came of it */
        return;
    }
    t->organ.ready = true;
    t->organ.own_tables = tg;
    t->organ.scratch.resize(TBF_FRAGMENT);
}

void initVibrato(void *, void *) {} /* the scanner's tables are part of the tone generator's */

/* frees the organ's run-time state and its wave tables on the device too: the CLAP host does
 * freeToneGenerator + allocTonegen + initToneGenerator on every tuning change
 * (src/clap.cpp:129-160) */
void freeToneGenerator(struct b_tonegen *t)
{
    if (!t)
        return;
    if (t->organ.ready && g.ctx) {
        tbf_gpu_instance_destroy(g.ctx, t->organ.id, 1);
        tbf_gpu_tonegen_tables_destroy(g.ctx, t->organ.own_tables);
    }
    delete t;
    --g.live;
}

void oscKeyOn(struct b_tonegen *t, short midiNote, short) { t->organ.push(TBF_EV_KEY_ON, midiNote); }
void oscKeyOff(struct b_tonegen *t, short midiNote, short) { t->organ.push(TBF_EV_KEY_OFF, midiNote); }

void setDrawBar(struct b_tonegen *t, int bus, unsigned int setting)
{
    t->organ.push(TBF_EV_DRAWBAR, bus, (int)setting);
}

void setDrawBars(void *inst, unsigned int manual, unsigned int setting[])
{
    if (manual > 2)
        return; /* the reference asserts */
    for (int i = 0; i < 9; ++i)
        setDrawBar(((b_instance *)inst)->synth, (int)manual * 9 + i, setting[i]);
}

void setPercussionEnabled(struct b_tonegen *t, int v) { t->organ.push(TBF_EV_PERC_ENABLE, v); }
void setPercussionVolume(struct b_tonegen *t, int v) { t->organ.push(TBF_EV_PERC_SOFT, v); }
void setPercussionFast(struct b_tonegen *t, int v) { t->organ.push(TBF_EV_PERC_FAST, v); }
void setPercussionFirst(struct b_tonegen *t, int v) { t->organ.push(TBF_EV_PERC_FIRST, v); }
void setVibratoUpper(struct b_tonegen *t, int v) { t->organ.push(TBF_EV_VIB_UPPER, v); }
void setVibratoLower(struct b_tonegen *t, int v) { t->organ.push(TBF_EV_VIB_LOWER, v); }
void setVibratoFromInt(void *t, int param) { ((b_tonegen *)t)->organ.push(TBF_EV_VIB_KNOB, param); }

void oscGenerateFragment(struct b_tonegen *t, float *buf, size_t)
{
    t->organ.render(1, TBF_RENDER_WHIRL_BYPASS, nullptr, buf, t->organ.scratch.data());
}

/* ---------------- Leslie ---------------- */

struct b_whirl *allocWhirl(void)
{
    ++g.live;
    return new b_whirl();
}

int whirlConfig(struct b_whirl *w, ConfigContext *cfg) { return store_cfg(w->cfg, "whirl.", cfg); }

void initWhirl(struct b_whirl *w, void *, double rate)
{
    tbf_ctx *ctx = context();
    if (!ctx)
        return;
    tbf_tonegen_spec ts;
    tbf_whirl_spec ws;
    tbf_tonegen_spec_default(&ts, rate);
    tbf_whirl_spec_default(&ws, rate);
    for (const auto &kv : w->cfg)
        tbf_spec_apply_cfg(&ts, &ws, kv.first.c_str(), kv.second.c_str());
    tbf_whirl_tables wt;
    int rc = tbf_build_whirl_tables(&ws, &wt);
    if (rc != TBF_OK) {
        fail(rc, "tbf_build_whirl_tables");
        return;
    }
    int wh = -1;
    rc = tbf_gpu_whirl_tables_create(ctx, &wt, &wh);
    tbf_free_whirl_tables(&wt);
    const int tg = rc == TBF_OK ? default_tonegen_tables(ctx) : -1;
    if (rc == TBF_OK && tg >= 0)
        rc = tbf_gpu_instance_create(ctx, tg, wh, 1, &w->organ.id);
    if (rc != TBF_OK || tg < 0) {
        fail(rc, "Leslie upload");
        if (wh >= 0)
            tbf_gpu_whirl_tables_destroy(ctx, wh);
        return;
    }
    w->organ.ready = true;
    w->organ.own_tables = wh;
}

void freeWhirl(struct b_whirl *w)
{
    if (!w)
        return;
    if (w->organ.ready && g.ctx) {
        tbf_gpu_instance_destroy(g.ctx, w->organ.id, 1);
        tbf_gpu_whirl_tables_destroy(g.ctx, w->organ.own_tables);
    }
    delete w;
    --g.live;
}

/* `signals` bit 1 (notify the MIDI layer) has no meaning here; bit 2 (track revSelect) is
 * always on, as in every call the hosts make (src/whirl.cpp:174-224) */
void useRevOption(struct b_whirl *w, int n, int) { w->organ.push(TBF_EV_REV_OPTION, n); }
void setRevSelect(struct b_whirl *w, int n) { w->organ.push(TBF_EV_REV_SELECT, n); }

void whirlProc(struct b_whirl *w, const float *in, float *outL, float *outR, size_t n)
{
    if (n == 0 || n % TBF_FRAGMENT)
        return fail(TBF_ERR_ARG, "whirlProc: length not a multiple of 128");
    w->organ.render((uint32_t)(n / TBF_FRAGMENT), TBF_RENDER_WHIRL_INPUT, in, outL, outR);
}

/* the horn and drum stems next to (or instead of) the mix; NULL outputs are skipped, as in the
 * reference (src/whirl.cpp:1191-1215, :1598-1612) */
void whirlProc2(struct b_whirl *w, const float *in, float *outL, float *outR, float *outHL, float *outHR,
                float *outDL, float *outDR, size_t n)
{
    if (n == 0 || n % TBF_FRAGMENT)
        return fail(TBF_ERR_ARG, "whirlProc2: length not a multiple of 128");
    float *stems[4] = {outHL, outHR, outDL, outDR};
    if (!outL || !outR) { /* the batch API always forms the mix: give it somewhere to go */
        w->organ.scratch.resize(2 * n);
        if (!outL)
            outL = w->organ.scratch.data();
        if (!outR)
            outR = w->organ.scratch.data() + n;
    }
    w->organ.render((uint32_t)(n / TBF_FRAGMENT), TBF_RENDER_WHIRL_INPUT, in, outL, outR, stems);
}

void whirlProc3(struct b_whirl *w, const float *in, float *outL, float *outR, float *, float *, size_t n)
{
    if (n == 0 || n % TBF_FRAGMENT)
        return fail(TBF_ERR_ARG, "whirlProc3: length not a multiple of 128");
    w->organ.render((uint32_t)(n / TBF_FRAGMENT), TBF_RENDER_WHIRL_INPUT | TBF_RENDER_WHIRLPROC3, in, outL,
                    outR);
}

/* ---------------- process-wide ---------------- */

int tbf_shim_status(void) { return g.status; }
const char *tbf_shim_last_error(void) { return g.error.c_str(); }

void tbf_shim_shutdown(void)
{
    if (g.ctx)
        tbf_gpu_destroy(g.ctx);
    g = Shim();
}

} // extern "C"
