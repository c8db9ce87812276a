/* tbf_kernels.cu -- sm_100a kernels for the tonewheel -> scanner -> Leslie path.
 *
 * One CTA of 128 threads owns one organ instance for a whole batch of
 * fragments.  Thread n is sample n of the 128-sample fragment for every stage
 * without a recurrence: the additive wheel sum, envelopes, percussion
 * high-pass, the Leslie tap geometry + 5-tap FIRs and the output mixes.  The
 * reference's two scatter-add delay structures (scanner ring, four Leslie delay
 * lines) are evaluated as ORDERED GATHERS, thread = destination slot, adding
 * the contributions in exactly the order the reference's sample loop does.
 * Only the true recurrences stay on single lanes, spread over different warps
 * so they overlap: four biquads, two rotor angles, the percussion / key
 * compression ramps and the control-event bookkeeping.
 * Per-instance scalars and the live window of every delay line sit in shared
 * memory for the whole launch.
 *
 * Numerics contract: bit-exact float/double arithmetic in the
 * reference's source operation order.  This file is compiled with
 * -fmad=false -ftz=false -prec-div=true -prec-sqrt=true; no FMA contraction,
 * no re-association, denormals kept.
 *
 * Reference functions restated here (never copied):
 *   oscKeyOn/oscKeyOff          src/tonegen.cpp:3096-3166
 *   oscGenerateFragment         src/tonegen.cpp:3218-3778
 *   vibratoProc                 src/vibrato.cpp:365-411
 *   whirlProc2 / whirlProc3     src/whirl.cpp:1191-1681
 *   setters                     src/tonegen.cpp:1635-1765, :2738-2750; src/vibrato.cpp:95-142;
 *                               src/whirl.cpp:174-236
 */
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "tbf_device.h"

#define DISPLC_MASK (TBF_WHIRL_DISPLC - 1)
#define GLOBAL_WBUF_MASK (TBF_WHIRL_BUF - 1)
#define VRING 256            /* shared-memory scanner ring: writes land <= 27 slots ahead */
#define VRING_MASK (VRING - 1)
#define VWINDOW (TBF_FRAGMENT + 100)

namespace {

/* One entry of the per-fragment "core program": the reference emits 1-2 CoreIns
 * per active wheel (src/tonegen.h:114-129); since every thread is one sample, the
 * wrap split collapses into a modular index and one record per wheel suffices. */
struct CoreRec {
    uint32_t wave_off;
    uint32_t len;
    uint32_t pos;
    int32_t env; /* -1: plain gain; 0..7 attack row; 8..15 release row */
    float gs, gv, gp;
    float ds, dv, dp;
};

struct __align__(16) Smem {
    InstScalars sc;
    CoreRec rec[TBF_NOF_WHEELS + 1];
    uint8_t removed[TBF_NOF_WHEELS + 3];
    uint16_t removed_wheel[TBF_NOF_WHEELS + 1];
    float vib_in[TBF_FRAGMENT], prc[TBF_FRAGMENT];
    float pg[TBF_FRAGMENT], kc[TBF_FRAGMENT], mono[TBF_FRAGMENT];
    /* scanner: per-sample contributions, inverse slot map, ring */
    float va[TBF_FRAGMENT], vb[TBF_FRAGMENT], vib_y[TBF_FRAGMENT];
    uint16_t vh[TBF_FRAGMENT];
    uint8_t vfirst[VRING];
    float vring[VRING];
    /* whirl */
    double hang[TBF_FRAGMENT], dang[TBF_FRAGMENT];
    float hx[3][8 + TBF_FRAGMENT]; /* horn stage inputs with 8 samples of history */
    float dx[3][1 + TBF_FRAGMENT]; /* drum stage inputs with 1 sample of history  */
    float leak[TBF_FRAGMENT];
    uint16_t tapn[12][TBF_FRAGMENT]; /* destination slot of the tap's main part (ring index) */
    float tapa[12][TBF_FRAGMENT], tapb[12][TBF_FRAGMENT];
    float lineout[4][TBF_FRAGMENT];  /* HL HR raw reads, DL DR after the drum biquad */
    unsigned long long csum;
    int n_rec, n_ins, recompute_routing, draw_bar_change, vib_on, perc_on;
    float key_comp_delta;
    float zs[4]; /* FILTER_C states z[0..3], NaN-scrubbed per fragment */
    /* followed by float wring[4][ring_size]: the Leslie delay lines */
};

__device__ __forceinline__ unsigned long long mix64(unsigned long long x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

/* x86 semantics of `(unsigned int)f` for the float->unsigned casts of
 * src/whirl.cpp:56-58: gcc converts through a 64-bit signed integer and keeps
 * the low 32 bits (CUDA's direct cast would saturate negatives to 0). */
__device__ __forceinline__ unsigned f2u_x86(float f) { return (unsigned)(long long)f; }

/* ------------------------------------------------------------------ */
/* control events: thread 0 only                                       */
/* ------------------------------------------------------------------ */

__device__ void key_off(InstScalars &sc, InstState *st, const DevTonegenTables *tg, int key)
{
    if (key < 0 || key >= TBF_MAX_KEYS)
        return;
    if (!st->activeKeys[key])
        return;
    st->activeKeys[key] = 0;
    if (key < 128)
        sc.upperKeyCount--;
    sc.keyDownCount--;
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0; e < e1; ++e) {
        const int w = tg->contrib_wheel[e];
        const int b = tg->contrib_bus[e];
        OscState *o = &st->osc[w];
        float *bl = &st->busLevel[w * TBF_NOF_BUSES + b];
        *bl = *bl - tg->contrib_level[e];
        o->refCount -= 1;
        if (o->refCount == 0)
            o->rflags = OR_REM;
        else
            o->rflags |= ORF_MODIFIED;
    }
}

__device__ void key_on(InstScalars &sc, InstState *st, const DevTonegenTables *tg, int key)
{
    if (key < 0 || key >= TBF_MAX_KEYS)
        return;
    if (st->activeKeys[key])
        key_off(sc, st, tg, key);
    st->activeKeys[key] = 1;
    if (key < 128)
        sc.upperKeyCount++;
    sc.keyDownCount++;
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0; e < e1; ++e) {
        const int w = tg->contrib_wheel[e];
        const int b = tg->contrib_bus[e];
        OscState *o = &st->osc[w];
        if (o->refCount == 0) {
            o->rflags = OR_ADD;
            if (o->aclPos == -1) {
                o->aclPos = sc.activeOscLEnd;
                st->activeOscList[sc.activeOscLEnd++] = w;
            }
        } else {
            o->rflags |= ORF_MODIFIED;
        }
        float *bl = &st->busLevel[w * TBF_NOF_BUSES + b];
        *bl = *bl + tg->contrib_level[e];
        o->refCount += 1;
    }
}

__device__ void perc_resets(InstScalars &sc, const DevTonegenTables *tg)
{
    if (sc.percIsFast)
        sc.percEnvGainDecay = sc.percIsSoft ? tg->perc_decay_fast_soft : tg->perc_decay_fast_norm;
    else
        sc.percEnvGainDecay = sc.percIsSoft ? tg->perc_decay_slow_soft : tg->perc_decay_slow_norm;
}

__device__ void use_rev_option(InstScalars &sc, const DevWhirlTables *wh, int n, int signals)
{
    const int i = n % 9;
    sc.hornTarget = wh->rev_horn_target[i];
    sc.drumTarget = wh->rev_drum_target[i];
    if (sc.hornIncr < sc.hornTarget)
        sc.hornAcDc = 1;
    else if (sc.hornTarget < sc.hornIncr)
        sc.hornAcDc = -1;
    if (sc.drumIncr < sc.drumTarget)
        sc.drumAcDc = 1;
    else if (sc.drumTarget < sc.drumIncr)
        sc.drumAcDc = -1;
    if (signals & 2) {
        const int hr = (n / 3) % 3;
        sc.revSelect = (hr == 2) ? 2 : (hr == 1) ? 0 : 1; /* WHIRL_FAST / SLOW / STOP */
    }
}

__device__ void apply_event(InstScalars &sc, InstState *st, const DevTonegenTables *tg,
                            const DevWhirlTables *wh, const tbf_event &ev)
{
    const int a = ev.a, b = ev.b;
    switch (ev.type) {
    case TBF_EV_KEY_ON:
        key_on(sc, st, tg, a);
        break;
    case TBF_EV_KEY_OFF:
        key_off(sc, st, tg, a);
        break;
    case TBF_EV_DRAWBAR:
        if (a < 0 || a >= TBF_NOF_BUSES || b < 0 || b >= 9) { /* reference asserts */
            sc.errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        sc.drawBarChange = 1;
        if (a == tg->perc_trigger_bus) {
            sc.percTrigRestore = b;
            if (sc.percEnabled)
                break;
        }
        sc.drawBarGain[a] = tg->draw_bar_level[a][b];
        break;
    case TBF_EV_PERC_ENABLE:
        if (a) {
            sc.newRouting |= RT_PERC;
            if (-1 < tg->perc_trigger_bus) {
                sc.drawBarGain[tg->perc_trigger_bus] = 0.0f;
                sc.drawBarChange = 1;
            }
        } else {
            sc.newRouting &= ~RT_PERC;
            if (-1 < tg->perc_trigger_bus) {
                sc.drawBarGain[tg->perc_trigger_bus] =
                    tg->draw_bar_level[tg->perc_trigger_bus][sc.percTrigRestore];
                sc.drawBarChange = 1;
            }
        }
        sc.percEnabled = a;
        break;
    case TBF_EV_PERC_SOFT:
        sc.percIsSoft = a;
        sc.percEnvGainReset = __fmul_rn(
            tg->perc_env_scaling, a ? tg->perc_env_gain_reset_soft : tg->perc_env_gain_reset_norm);
        sc.percDrawbarGain = a ? tg->perc_drawbar_soft_gain : tg->perc_drawbar_normal_gain;
        perc_resets(sc, tg);
        break;
    case TBF_EV_PERC_FAST:
        sc.percIsFast = a;
        perc_resets(sc, tg);
        break;
    case TBF_EV_PERC_FIRST:
        sc.percSendBus = a ? tg->perc_send_bus_a : tg->perc_send_bus_b;
        break;
    case TBF_EV_VIB_UPPER:
        if (a)
            sc.newRouting |= RT_UPPRVIB;
        else
            sc.newRouting &= ~RT_UPPRVIB;
        break;
    case TBF_EV_VIB_LOWER:
        if (a)
            sc.newRouting |= RT_LOWRVIB;
        else
            sc.newRouting &= ~RT_LOWRVIB;
        break;
    case TBF_EV_VIB_KNOB:
        if (a >= 0 && a <= 5) { /* V1 C1 V2 C2 V3 C3 */
            sc.vibTable = a >> 1;
            sc.vibMixed = (a & 1) ? 0x80 : 0;
        }
        break;
    case TBF_EV_REV_OPTION:
        if (a < 0) {
            sc.errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        use_rev_option(sc, wh, a, 2);
        break;
    case TBF_EV_REV_SELECT:
        if (a < 0) {
            sc.errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        sc.revSelect = a % 3;
        use_rev_option(sc, wh, wh->revselects[sc.revSelect], 1);
        break;
    case TBF_EV_SWELL:
        sc.swellPedalGain =
            (float)(((double)tg->output_level_trim * (double)(unsigned char)a) / 127.0);
        break;
    default:
        sc.errorFlags |= TBF_IFLAG_BAD_EVENT;
        break;
    }
}

/* ------------------------------------------------------------------ */
/* tone generator: active-list pass (src/tonegen.cpp:3333-3594)        */
/* ------------------------------------------------------------------ */

/* Entry i of the active list is independent of every other entry (it touches
 * only its own wheel's state), so thread i handles entries i, i+128, ... */
__device__ void active_list_pass(InstState *st, const DevTonegenTables *tg, Smem &sm, int tid)
{
    InstScalars &sc = sm.sc;
    const int n_act = sc.activeOscLEnd;
    const uint32_t routing = sc.oldRouting; /* already updated by thread 0 */
    const bool recompute_routing = sm.recompute_routing != 0;
    const bool draw_bar_change = sm.draw_bar_change != 0;
    int my_ins = 0;
    int any_removed = 0;

    for (int i = tid; i < n_act; i += TBF_THREADS) {
        const int w = st->activeOscList[i];
        OscState *o = &st->osc[w];
        const uint32_t rflags = o->rflags;
        CoreRec r;
        r.wave_off = tg->wave_offset[w];
        r.len = tg->wave_length[w];
        r.pos = o->pos;

        if (rflags & ORF_REMOVED) {
            sm.removed[i] = 1;
            any_removed = 1;
            r.env = 8 + (i & 7);
            r.gs = o->sumSwell;
            r.gp = o->sumPercn;
            r.gv = o->sumScanr;
            r.ds = __fsub_rn(0.0f, r.gs);
            r.dp = __fsub_rn(0.0f, r.gp);
            r.dv = __fsub_rn(0.0f, r.gv);
        } else {
            sm.removed[i] = 0;
            bool reroute = false;
            if (rflags & ORF_ADDED) {
                r.gs = r.gp = r.gv = 0.0f;
            } else {
                r.gs = o->sumSwell;
                r.gp = o->sumPercn;
                r.gv = o->sumScanr;
            }
            if ((rflags & ORF_MODIFIED) || draw_bar_change) {
                const float *bl = &st->busLevel[w * TBF_NOF_BUSES];
                float sum = 0.0f;
                for (int d = 0; d < 9; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], sc.drawBarGain[d]));
                o->sumUpper = sum;
                sum = 0.0f;
                for (int d = 9; d < 18; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], sc.drawBarGain[d]));
                o->sumLower = sum;
                sum = 0.0f;
                for (int d = 18; d < 27; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], sc.drawBarGain[d]));
                o->sumPedal = sum;
                reroute = true;
            }
            if (reroute || recompute_routing) {
                float percn = 0.0f;
                if (routing & RT_PERC)
                    percn = st->busLevel[w * TBF_NOF_BUSES + sc.percSendBus];
                float scanr = 0.0f;
                float swell = o->sumPedal;
                if (routing & RT_UPPRVIB)
                    scanr = __fadd_rn(scanr, o->sumUpper);
                else
                    swell = __fadd_rn(swell, o->sumUpper);
                if (routing & RT_LOWRVIB)
                    scanr = __fadd_rn(scanr, o->sumLower);
                else
                    swell = __fadd_rn(swell, o->sumLower);
                o->sumPercn = percn;
                o->sumScanr = scanr;
                o->sumSwell = swell;
            }
            if (rflags & OR_ADD) {
                r.env = i & 7;
                r.ds = __fsub_rn(o->sumSwell, r.gs);
                r.dp = __fsub_rn(o->sumPercn, r.gp);
                r.dv = __fsub_rn(o->sumScanr, r.gv);
            } else {
                r.env = -1;
                r.ds = r.dp = r.dv = 0.0f;
            }
        }
        /* advance the read position; wrap test is `length < pos + 128`
         * (src/tonegen.cpp:3376, :3524), so pos may rest at == length */
        if (r.len < r.pos + TBF_FRAGMENT) {
            o->pos = TBF_FRAGMENT - (r.len - r.pos);
            my_ins += 2;
        } else {
            o->pos = r.pos + TBF_FRAGMENT;
            my_ins += 1;
        }
        o->rflags = 0;
        sm.rec[i] = r;
    }
    if (my_ins)
        atomicAdd(&sm.n_ins, my_ins);
    any_removed = __syncthreads_or(any_removed);

    if (tid == 0) {
        sm.n_rec = n_act;
        sc.statFragments += 1;
        sc.statActiveWheels += (unsigned long long)n_act;
        sc.drawBarChange = 0;
        if (sm.n_ins > 256)
            sc.errorFlags |= TBF_IFLAG_COREPGM_OVERFLOW;
        if (any_removed) {
            /* removal list in ascending active-list index, swap-with-last
             * (src/tonegen.cpp:3576-3594); changes the list order and with it
             * next fragment's summation order and envelope rows */
            int l_end = n_act;
            /* victims must be identified by wheel number BEFORE any swap moves
             * entries around: the reference collects removedList[] first */
            for (int i = 0; i < n_act; ++i)
                if (sm.removed[i])
                    sm.removed_wheel[i] = (uint16_t)st->activeOscList[i];
            for (int i = 0; i < n_act; ++i) {
                if (!sm.removed[i])
                    continue;
                const int vic = sm.removed_wheel[i];
                const int actidx = st->osc[vic].aclPos;
                st->osc[vic].aclPos = -1;
                l_end--;
                if (0 < l_end) {
                    const int mov = st->activeOscList[l_end];
                    if (mov != vic) {
                        st->activeOscList[actidx] = mov;
                        st->osc[mov].aclPos = actidx;
                    }
                }
            }
            sc.activeOscLEnd = l_end;
        }
    }
}

/* ------------------------------------------------------------------ */
/* tone generator: core interpreter, thread = sample                   */
/* (src/tonegen.cpp:3607-3687)                                         */
/* ------------------------------------------------------------------ */
__device__ __forceinline__ void core_interpreter(const DevTonegenTables *tg, const Smem &sm, int n,
                                                 float &swl, float &vib, float &prc)
{
    const int n_rec = sm.n_rec;
    const float *__restrict__ wave = tg->wave_data;
    float s = 0.0f, v = 0.0f, p = 0.0f;
    for (int i = 0; i < n_rec; ++i) {
        const CoreRec &r = sm.rec[i];
        uint32_t idx = r.pos + (uint32_t)n;
        if (idx >= r.len)
            idx -= r.len;
        const float x = __ldg(wave + r.wave_off + idx);
        float cs, cv, cp;
        if (r.env >= 0) {
            const float e = __ldg(&tg->env[r.env][n]);
            cs = __fmul_rn(x, __fadd_rn(r.gs, __fmul_rn(e, r.ds)));
            cv = __fmul_rn(x, __fadd_rn(r.gv, __fmul_rn(e, r.dv)));
            cp = __fmul_rn(x, __fadd_rn(r.gp, __fmul_rn(e, r.dp)));
        } else {
            cs = __fmul_rn(x, r.gs);
            cv = __fmul_rn(x, r.gv);
            cp = __fmul_rn(x, r.gp);
        }
        if (i == 0) { /* first instruction of a fragment copies, the rest add */
            s = cs;
            v = cv;
            p = cp;
        } else {
            s = __fadd_rn(s, cs);
            v = __fadd_rn(v, cv);
            p = __fadd_rn(p, cp);
        }
    }
    swl = s;
    vib = v;
    prc = p;
}

/* ------------------------------------------------------------------ */
/* scanner vibrato (src/vibrato.cpp:365-411)                            */
/* ------------------------------------------------------------------ */

/* Per-sample part, thread = sample m.  Everything integer is closed form in m:
 * stator(m) = stator0 + m*inc, outPos(m) = outPos0 + m.  Sample m adds (x - g) to
 * slot h(m) and g to slot h(m)+1.  Returns non-zero when the gather's assumptions
 * (slots advance monotonically, stay inside the shared ring) do not hold. */
__device__ __forceinline__ int vibrato_contrib(const InstScalars &sc, const DevTonegenTables *tg,
                                               Smem &sm, int m, float x)
{
    const float fnorm = (float)(1.0 / 65536.0);
    const uint32_t *tab = tg->vib_offset_table[sc.vibTable];
    const uint32_t inc = tg->vib_stator_increment;
    const uint32_t stator = (sc.vibStator + (uint32_t)m * inc) & 0x07FFFFFFu;
    const uint32_t op = (sc.vibOutPos + (uint32_t)m) & 0x3FFu;
    const uint32_t j = ((op << 16) + __ldg(&tab[stator >> 16])) & 0x03FFFFFFu;
    const uint32_t h = j >> 16;
    const float f = __fmul_rn(fnorm, (float)(j & 0xFFFFu));
    const float g = __fmul_rn(f, x);
    sm.va[m] = __fsub_rn(x, g);
    sm.vb[m] = g;
    sm.vh[m] = (uint16_t)h;
    const uint32_t rel = (h - sc.vibOutPos) & 0x3FFu; /* slots ahead of the fragment's first read */
    int bad = (rel < (uint32_t)m) || (rel + 1 >= (uint32_t)VWINDOW);
    uint32_t hprev = 0xFFFFFFFFu;
    if (m > 0) {
        const uint32_t sp = (sc.vibStator + (uint32_t)(m - 1) * inc) & 0x07FFFFFFu;
        const uint32_t opp = (sc.vibOutPos + (uint32_t)(m - 1)) & 0x3FFu;
        hprev = (((opp << 16) + __ldg(&tab[sp >> 16])) & 0x03FFFFFFu) >> 16;
        bad |= ((h - hprev) & 0x3FFu) >= 512u; /* went backwards */
    }
    if (!bad && h != hprev)
        sm.vfirst[rel] = (uint8_t)m;
    return bad;
}

/* Gather, thread = slot s ahead of the fragment's first read position.  The slot
 * collects, in sample order, the g-parts of the samples that hit slot s-1 and then
 * the (x-g)-parts of the samples that hit s -- the order the reference's loop adds
 * them in, because h(m) never decreases.  Slots s < 128 are read (and cleared) by
 * sample s after all of its contributors (they are >= 1 sample older). */
__device__ __forceinline__ float vibrato_gather_slot(const InstScalars &sc, Smem &sm, int s)
{
    const uint32_t S = (sc.vibOutPos + (uint32_t)s) & 0x3FFu;
    float v = sm.vring[S & VRING_MASK];
    if (s > 0) {
        const uint32_t Sm1 = (S - 1) & 0x3FFu;
        for (int m = sm.vfirst[s - 1]; m < TBF_FRAGMENT && sm.vh[m] == Sm1; ++m)
            v = __fadd_rn(v, sm.vb[m]);
    }
    for (int m = sm.vfirst[s]; m < TBF_FRAGMENT && sm.vh[m] == S; ++m)
        v = __fadd_rn(v, sm.va[m]);
    return v;
}

/* the reference's loop as written, one lane: fallback when the gather's
 * assumptions fail (pathological offset tables) */
__device__ void vibrato_serial(InstScalars &sc, const DevTonegenTables *tg, Smem &sm)
{
    const float fnorm = (float)(1.0 / 65536.0);
    const float mixnorm = (float)0.7071067811865475;
    const uint32_t *tab = tg->vib_offset_table[sc.vibTable];
    const uint32_t inc = tg->vib_stator_increment;
    const int mixed = sc.vibMixed;
    uint32_t stator = sc.vibStator, out_pos = sc.vibOutPos;
    float *buf = sm.vring;
    for (int i = 0; i < TBF_FRAGMENT; ++i) {
        const float x = sm.vib_in[i];
        const uint32_t j = ((out_pos << 16) + __ldg(&tab[stator >> 16])) & 0x03FFFFFFu;
        const uint32_t h = j >> 16;
        const uint32_t k = (h + 1) & 0x3FFu;
        const float f = __fmul_rn(fnorm, (float)(j & 0xFFFFu));
        const float g = __fmul_rn(f, x);
        buf[h & VRING_MASK] = __fadd_rn(buf[h & VRING_MASK], __fsub_rn(x, g));
        buf[k & VRING_MASK] = __fadd_rn(buf[k & VRING_MASK], g);
        float y;
        if (mixed)
            y = __fmul_rn(__fadd_rn(x, buf[out_pos & VRING_MASK]), mixnorm);
        else
            y = buf[out_pos & VRING_MASK];
        sm.vib_y[i] = y;
        buf[out_pos & VRING_MASK] = 0.0f;
        out_pos = (out_pos + 1) & 0x3FFu;
        stator = (stator + inc) & 0x07FFFFFFu;
    }
}

/* percussion-gain decay and key-compression chase: two scalar recurrences
 * carried across fragments (src/tonegen.cpp:3733-3771), one lane */
__device__ void gain_ramps(InstScalars &sc, Smem &sm, bool perc_on)
{
    float kc = sc.keyCompLevel;
    const float delta = sm.key_comp_delta;
    if (perc_on) {
        float pg = sc.percEnvGain;
        const float decay = sc.percEnvGainDecay;
#pragma unroll 8
        for (int i = 0; i < TBF_FRAGMENT; ++i) {
            sm.pg[i] = pg;
            sm.kc[i] = kc;
            pg = __fmul_rn(pg, decay);
            kc = __fadd_rn(kc, delta);
        }
        sc.percEnvGain = pg;
    } else {
#pragma unroll 8
        for (int i = 0; i < TBF_FRAGMENT; ++i) {
            sm.kc[i] = kc;
            kc = __fadd_rn(kc, delta);
        }
    }
    sc.keyCompLevel = kc;
}

/* ------------------------------------------------------------------ */
/* whirl: per-call speed ramp and brake logic of ONE rotor, double      */
/* precision, one lane (src/whirl.cpp:1219-1374; the horn and drum      */
/* halves there differ only in the constants passed here).              */
/* Returns non-zero when the brake is engaged (incr is zeroed after the */
/* sample loop, src/whirl.cpp:1635-1636).                               */
/* ------------------------------------------------------------------ */
struct RotorConst {
    float hardstop_rpm;  /* 10 horn, 8 drum      */
    float brake_limit_rpm; /* 60 horn, 100 drum  */
    int is_horn;
};

__device__ int rotor_ramp(const DevWhirlTables *wh, const RotorConst rc, int *acdc, double *incr,
                          double *angle, const double target, const double brake_pos,
                          const double oml_acc, const double oml_dec)
{
    const double SR = wh->sample_rate;
    const float nsmp = (float)TBF_FRAGMENT;
    const double target_pos =
        rc.is_horn ? fmod(__dsub_rn(1.25, brake_pos), 1.0) : fmod(__dadd_rn(brake_pos, .75), 1.0);

    if (*acdc) {
        int flywheel = 0;
        const float hardstop = (float)((double)rc.hardstop_rpm / ((double)60.f * SR));
        if (brake_pos > 0 && target == 0 && *incr > 0 && *incr < (double)hardstop) {
            if (fabs(__dsub_rn(*angle, target_pos)) < (2.0 / TBF_WHIRL_DISPLC)) {
                *angle = target_pos;
                *incr = 0;
            } else {
                const float minspeed = (float)((double)3.f / ((double)60.f * SR));
                const float diffinc = (float)(
                    fmod(__dsub_rn(__dadd_rn(1., target_pos), *angle), 1.0) / (double)nsmp);
                if (*incr > (double)diffinc)
                    *incr = diffinc;
                else if (*incr < (double)minspeed)
                    *incr = minspeed;
                flywheel = 1;
            }
        }
        if (!flywheel) {
            const double oml = (*acdc > 0) ? oml_acc : oml_dec;
            *incr = __dadd_rn(*incr, __dmul_rn(oml, __dsub_rn(target, *incr)));
        }
        if (fabs(__dsub_rn(target, *incr)) < (.05 / ((double)60.f * SR))) {
            *acdc = 0;
            *incr = target;
        }
    }

    int brake = 0;
    if (brake_pos > 0) {
        if (!*acdc && *incr == 0 && *angle != target_pos) {
            brake = 1;
            if (fabs(__dsub_rn(*angle, target_pos)) < (2.0 / TBF_WHIRL_DISPLC)) {
                *angle = target_pos;
            } else {
                const float limit = (float)((double)rc.brake_limit_rpm / (60. * SR));
                *incr = fmod(__dsub_rn(__dadd_rn(1., target_pos), *angle), 1.0) / (double)nsmp;
                if (*incr > (double)limit)
                    *incr = limit;
            }
        }
    }
    return brake;
}

/* rotor angle recurrence, angle = fmod(angle + incr, 1.0) in double
 * (src/whirl.cpp:1618-1619); for 0 <= t < 2 fmod(t,1) is t or t-1 exactly */
__device__ __forceinline__ double angle_step(double a, double incr)
{
    const double t = __dadd_rn(a, incr);
    if (t >= 0.0 && t < 1.0)
        return t;
    if (t >= 1.0 && t < 2.0)
        return __dsub_rn(t, 1.0);
    return fmod(t, 1.0);
}

__device__ void angle_lane(double a0, double incr, double *out, double *final_angle)
{
    double a = a0;
    if (a0 >= 0.0 && a0 < 1.0 && incr >= 0.0 && incr < 1.0) {
        /* every sum stays in [0, 2): fmod(t, 1.0) is t or t - 1, both exact */
#pragma unroll 8
        for (int i = 0; i < TBF_FRAGMENT; ++i) {
            out[i] = a;
            const double t = __dadd_rn(a, incr);
            a = (t >= 1.0) ? __dsub_rn(t, 1.0) : t;
        }
    } else {
        for (int i = 0; i < TBF_FRAGMENT; ++i) {
            out[i] = a;
            a = angle_step(a, incr);
        }
    }
    *final_angle = a;
}

/* speed ramp + angle sequence of one rotor; the horn and the drum lane run this in
 * lockstep on two lanes of one warp */
__device__ void rotor_lane(InstScalars &sc, const DevWhirlTables *wh, Smem &sm, bool horn)
{
    RotorConst rc;
    rc.hardstop_rpm = horn ? 10.f : 8.f;
    rc.brake_limit_rpm = horn ? 60.f : 100.f;
    rc.is_horn = horn ? 1 : 0;
    int *p_acdc = horn ? &sc.hornAcDc : &sc.drumAcDc;
    double *p_incr = horn ? &sc.hornIncr : &sc.drumIncr;
    double *p_ang = horn ? &sc.hornAngle : &sc.drumAngle;
    int acdc = *p_acdc;
    double incr = *p_incr, ang = *p_ang;
    const int brake = rotor_ramp(wh, rc, &acdc, &incr, &ang, horn ? sc.hornTarget : sc.drumTarget,
                                 horn ? wh->hn_brake_pos : wh->dr_brake_pos,
                                 wh->one_minus_l[horn ? 0 : 2], wh->one_minus_l[horn ? 1 : 3]);
    double fin;
    angle_lane(ang, incr, horn ? sm.hang : sm.dang, &fin);
    *p_acdc = acdc;
    *p_incr = brake ? 0.0 : incr;
    *p_ang = fin;
}

/* DF-II biquad, float, source operation order (EQ_IIR, src/whirl.cpp:1479-1485) */
__device__ __forceinline__ float eq_iir(const float *c, float &z0, float &z1, float x)
{
    const float t = __fsub_rn(__fsub_rn(x, __fmul_rn(c[0], z0)), __fmul_rn(c[1], z1));
    const float y = __fadd_rn(__fadd_rn(__fmul_rn(t, c[2]), __fmul_rn(c[3], z0)), __fmul_rn(c[4], z1));
    z1 = z0;
    z0 = t;
    return y;
}


/* horn input filters A and B in series (src/whirl.cpp:1513-1525), software-pipelined
 * over lanes 0 and 1 of warp 0: in iteration it lane 0 runs filter A on sample it while
 * lane 1 runs filter B on filter A's output of sample it-1 (handed over by shuffle), so
 * both recurrences share one instruction stream.  Call with tid 0 and 1 only. */
__device__ void horn_filter_lanes(InstScalars &sc, const DevWhirlTables *wh, Smem &sm, int lane)
{
    float c[5];
#pragma unroll
    for (int k = 0; k < 5; ++k)
        c[k] = wh->iir[lane][k];
    float z0 = sc.iirz[lane][0], z1 = sc.iirz[lane][1];
    const float leakage = wh->leakage;
    float handed = 0.0f; /* lane 1: filter A's output of the previous iteration */
#pragma unroll 4
    for (int it = 0; it <= TBF_FRAGMENT; ++it) {
        float in = handed;
        if (lane == 0 && it < TBF_FRAGMENT) {
            in = (float)__dadd_rn((double)sm.mono[it], 1e-14); /* DENORMAL_HACK */
            sm.dx[0][1 + it] = in;
        }
        const bool active = lane == 0 ? (it < TBF_FRAGMENT) : (it > 0);
        float y = 0.0f;
        if (active)
            y = eq_iir(c, z0, z1, in);
        handed = __shfl_sync(0x3u, y, 0);
        if (lane == 1 && it > 0) {
            sm.hx[0][8 + it - 1] = y;
            sm.leak[it - 1] = __fmul_rn(y, leakage);
        }
    }
    /* NaN scrub after the loop (EQ_IIR_NAN, src/whirl.cpp:1622-1625) */
    sc.iirz[lane][0] = isnan(z0) ? 0.0f : z0;
    sc.iirz[lane][1] = isnan(z1) ? 0.0f : z1;
}

/* drum filter of one channel on the values read from its delay line, one lane */
__device__ void drum_filter_lane(InstScalars &sc, const DevWhirlTables *wh, Smem &sm, int line)
{
    float c[5];
#pragma unroll
    for (int k = 0; k < 5; ++k)
        c[k] = wh->iir[line][k];
    float z0 = sc.iirz[line][0], z1 = sc.iirz[line][1];
    float *v = sm.lineout[line];
#pragma unroll 8
    for (int i = 0; i < TBF_FRAGMENT; ++i)
        v[i] = eq_iir(c, z0, z1, v[i]);
    sc.iirz[line][0] = isnan(z0) ? 0.0f : z0;
    sc.iirz[line][1] = isnan(z1) ? 0.0f : z1;
}

/* FILTER_C(0.4, 0.4): the literals are double, so the 2-tap FIR is evaluated in
 * double and rounded to float once (src/whirl.cpp:1472-1477) */
__device__ __forceinline__ float filter_c(float x, float z)
{
    return (float)__dadd_rn(__dmul_rn(0.4, (double)x), __dmul_rn(0.4, (double)z));
}

/* ------------------------------------------------------------------ */
/* whirl: tap geometry, thread = sample (HN_MOTION / DR_MOTION,         */
/* src/whirl.cpp:1432-1469).  Writes the slot and the two contributions */
/* of every tap to shared memory.                                       */
/* ------------------------------------------------------------------ */
__device__ __forceinline__ void tap_position(double ang_grd, int phase, const float *__restrict__ dsp,
                                             float spacing, uint32_t outpos, float &frac,
                                             uint32_t &slot, uint32_t &kround)
{
    const float h1 = (float)__dadd_rn(__dmul_rn(ang_grd, (double)TBF_WHIRL_DISPLC), (double)phase);
    const float hd = __fsub_rn(h1, truncf(h1)); /* fmodf(h1, 1.f), exact */
    const uint32_t hl = f2u_x86(floorf(h1)) & DISPLC_MASK;
    const uint32_t hh = (hl + 1) & DISPLC_MASK;
    const float intp =
        __fadd_rn(__fmul_rn(__ldg(dsp + hl), __fsub_rn(1.f, hd)), __fmul_rn(hd, __ldg(dsp + hh)));
    kround = f2u_x86(roundf(h1)) & DISPLC_MASK;
    const float t = __fadd_rn(__fadd_rn(spacing, intp), (float)outpos);
    const float r = floorf(t);
    frac = __fsub_rn(t, r);
    slot = f2u_x86(r) & GLOBAL_WBUF_MASK;
}

__device__ void whirl_taps(const DevWhirlTables *wh, Smem &sm, int i, uint32_t outpos0)
{
    const uint32_t outpos = (outpos0 + (uint32_t)i) & GLOBAL_WBUF_MASK;
    const double hang = sm.hang[i], dang = sm.dang[i];
#pragma unroll
    for (int p = 0; p < 6; ++p) {
        /* taps 0,3,4 read the forward displacement table with the backward FIR
         * bank, taps 1,2,5 the other way round (src/whirl.cpp:1540-1563) */
        const bool fwd = (p == 0 || p == 3 || p == 4);
        const float *dsp = fwd ? wh->hn_fwd_displ : wh->hn_bwd_displ;
        const float *bw = fwd ? wh->bbw : wh->bfw;
        const double ang = __dadd_rn(hang, (p & 1) ? wh->bw_ang : wh->fw_ang);
        float frac;
        uint32_t slot, k;
        tap_position(ang, wh->horn_phase[p], dsp, wh->horn_spacing[p], outpos, frac, slot, k);
        const float *hx = &sm.hx[p >> 1][8 + i];
        const float *b = bw + 5 * k;
        float xa = __fmul_rn(__ldg(b + 0), hx[0]);
        xa = __fadd_rn(xa, __fmul_rn(__ldg(b + 1), hx[-1]));
        xa = __fadd_rn(xa, __fmul_rn(__ldg(b + 2), hx[-2]));
        xa = __fadd_rn(xa, __fmul_rn(__ldg(b + 3), hx[-3]));
        xa = __fadd_rn(xa, __fmul_rn(__ldg(b + 4), hx[-4]));
        const float q = __fmul_rn(xa, frac);
        sm.tapn[p][i] = (uint16_t)slot;
        sm.tapa[p][i] = __fsub_rn(xa, q);
        sm.tapb[p][i] = q;
    }
#pragma unroll
    for (int p = 0; p < 6; ++p) {
        const bool fwd = (p == 0 || p == 3 || p == 4);
        const float *dsp = fwd ? wh->dr_fwd_displ : wh->dr_bwd_displ;
        float frac;
        uint32_t slot, k;
        tap_position(dang, wh->drum_phase[p], dsp, wh->drum_spacing[p], outpos, frac, slot, k);
        const float x = sm.dx[p >> 1][1 + i];
        const float q = __fmul_rn(x, frac);
        sm.tapn[6 + p][i] = (uint16_t)slot;
        sm.tapa[6 + p][i] = __fsub_rn(x, q);
        sm.tapb[6 + p][i] = q;
    }
}

/* What the parallel delay-line update relies on, checked per fragment: per tap the
 * destination slots advance monotonically with the sample index and stay inside
 * [reader, reader + window). */
__device__ __forceinline__ int whirl_check_taps(const DevWhirlTables *wh, const Smem &sm, int i,
                                                uint32_t outpos0)
{
    int bad = 0;
    const uint32_t window = (uint32_t)wh->window;
#pragma unroll
    for (int p = 0; p < 12; ++p) {
        const uint32_t n = sm.tapn[p][i];
        const uint32_t rel = (n - outpos0) & GLOBAL_WBUF_MASK;
        bad |= (rel < (uint32_t)i) | (rel + 1 >= window);
        if (i > 0)
            bad |= ((n - (uint32_t)sm.tapn[p][i - 1]) & GLOBAL_WBUF_MASK) >= (TBF_WHIRL_BUF / 2);
    }
    return bad;
}

/* Ordered scatter of ONE tap, thread = sample.  Because the tap's slot sequence n[i]
 * never decreases, the samples hitting a slot form a contiguous run, and the
 * reference's loop adds to slot S first the q-parts of the run that hit S-1, then the
 * (x-q)-parts of the run that hits S.  The first sample of each run applies exactly
 * that sequence to its slot; a slot that only receives q-parts (the run after it
 * skips it, or the fragment ends) is completed by the last sample of the run before.
 * Distinct threads touch distinct slots, so one pass needs no atomics; the three taps
 * of a line are applied in three passes, farthest delay first (see tap_order). */
__device__ __forceinline__ void whirl_tap_pass(const Smem &sm, float *ring, uint32_t rmask, int p, int i)
{
    const uint16_t *n = sm.tapn[p];
    const float *a = sm.tapa[p];
    const float *b = sm.tapb[p];
    const uint32_t ni = n[i];
    const uint32_t prev = i ? (uint32_t)n[i - 1] : 0xFFFFFFFFu;
    const uint32_t next = (i < TBF_FRAGMENT - 1) ? (uint32_t)n[i + 1] : 0xFFFFFFFFu;
    if (ni != prev) { /* first sample of the run hitting slot ni */
        float v = ring[ni & rmask];
        const uint32_t sm1 = (ni - 1) & GLOBAL_WBUF_MASK;
        int j0 = i;
        while (j0 > 0 && n[j0 - 1] == sm1)
            --j0;
        for (int j = j0; j < i; ++j)
            v = __fadd_rn(v, b[j]);
        for (int m = i; m < TBF_FRAGMENT && n[m] == ni; ++m)
            v = __fadd_rn(v, a[m]);
        ring[ni & rmask] = v;
    }
    if (ni != next) { /* last sample of its run */
        const uint32_t sp1 = (ni + 1) & GLOBAL_WBUF_MASK;
        if (next != sp1) { /* nobody hits ni+1 as a main slot in this fragment */
            float v = ring[sp1 & rmask];
            int j0 = i;
            while (j0 > 0 && n[j0 - 1] == ni)
                --j0;
            for (int j = j0; j <= i; ++j)
                v = __fadd_rn(v, b[j]);
            ring[sp1 & rmask] = v;
        }
    }
}

/* The reference's loop as written, one lane per delay line (0 HL, 1 HR, 2 DL,
 * 3 DR): fallback for geometries / rotor speeds the gather does not cover. */
__device__ void scatter_lane(const Smem &sm, float *buf, uint32_t rmask, float *lineout, int line,
                             uint32_t outpos0)
{
    const int tap0 = (line < 2) ? (line & 1) : 6 + (line & 1);
    for (int i = 0; i < TBF_FRAGMENT; ++i) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const int p = tap0 + 2 * s;
            const uint32_t n = sm.tapn[p][i];
            buf[n & rmask] = __fadd_rn(buf[n & rmask], sm.tapa[p][i]);
            const uint32_t n1 = (n + 1) & GLOBAL_WBUF_MASK;
            buf[n1 & rmask] = __fadd_rn(buf[n1 & rmask], sm.tapb[p][i]);
        }
        const uint32_t op = (outpos0 + (uint32_t)i) & GLOBAL_WBUF_MASK;
        lineout[i] = buf[op & rmask];
        buf[op & rmask] = 0.0f;
    }
}

/* ------------------------------------------------------------------ */
/* the batched render kernel: grid = instances, block = 128 threads     */
/* ------------------------------------------------------------------ */
__global__ void __launch_bounds__(TBF_THREADS)
tbf_render_kernel(const RenderParams p, const uint32_t ring_size)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
    float *wring = reinterpret_cast<float *>(smem_raw + sizeof(Smem));
    const uint32_t rmask = ring_size - 1;
    const int tid = threadIdx.x;
    const uint32_t slot = blockIdx.x;
    const uint32_t id = p.ids ? p.ids[slot] : slot;
    InstState *st = p.states + id;
    InstScalars &sc = sm.sc;

    /* ---- stage scalars, delay-line windows and FIR histories in shared memory ---- */
    for (int i = tid; i < (int)(sizeof(InstScalars) / 4); i += TBF_THREADS)
        reinterpret_cast<uint32_t *>(&sm.sc)[i] = reinterpret_cast<const uint32_t *>(&st->sc)[i];
    for (int i = tid; i < VRING; i += TBF_THREADS)
        sm.vfirst[i] = 0;
    if (tid == 0)
        sm.csum = 0ull;
    __syncthreads();
    const DevTonegenTables *tg = p.tg + sc.tgTables;
    const DevWhirlTables *wh = p.wh + sc.whTables;
    for (int s = tid; s < VRING; s += TBF_THREADS) {
        const uint32_t g = (sc.vibOutPos + (uint32_t)s) & 0x3FFu;
        sm.vring[g & VRING_MASK] = st->vibBuffer[g];
    }
    for (int line = 0; line < 4; ++line)
        for (uint32_t s = tid; s < ring_size; s += TBF_THREADS) {
            const uint32_t g = (sc.outpos + s) & GLOBAL_WBUF_MASK;
            wring[line * ring_size + (g & rmask)] = st->wbuf[line][g];
        }
    if (tid < 24) {
        const int s = tid >> 3, k = tid & 7;
        sm.hx[s][7 - k] = sc.adx[s][(sc.adi[s] + k) & 7];
    }
    if (tid < 4)
        sm.zs[tid] = sc.z[tid];

    const bool whirl_off = (p.flags & TBF_RENDER_WHIRL_BYPASS) != 0;
    const bool whirl3 = (p.flags & TBF_RENDER_WHIRLPROC3) != 0;
    const bool whirl_run = !whirl_off && !wh->bypass;
    const bool gather_ok = wh->gather_ok != 0;
    const int window = wh->window;
    const bool store = !(p.flags & TBF_RENDER_NO_OUTPUT);
    float *out_l = p.out_l + (size_t)slot * p.out_stride;
    float *out_r = p.out_r + (size_t)slot * p.out_stride;
    uint32_t outpos0 = sc.outpos;
    uint64_t ev_cur = p.ev_cursor[slot];
    const uint64_t ev_end = p.ev_start[slot + 1];
    unsigned long long csum = 0ull;
    __syncthreads();

    for (uint32_t f = 0; f < p.n_fragments; ++f) {
        /* -- P0: control events that precede this fragment, routing, key compression -- */
        if (tid == 0) {
            while (ev_cur < ev_end && p.events[ev_cur].fragment <= p.frag_base + f) {
                apply_event(sc, st, tg, wh, p.events[ev_cur]);
                ++ev_cur;
            }
            int kdc = sc.keyDownCount;
            if (kdc > 127 || kdc < 0) { /* reference reads past keyCompTable[128] */
                sc.errorFlags |= TBF_IFLAG_KEYCOMP_OOB;
                kdc = kdc < 0 ? 0 : 127;
            }
            const float key_comp = tg->key_comp_table[kdc];
            /* (keyComp - level) / 128.f: dividing by 2^7 and multiplying by 2^-7 round identically */
            sm.key_comp_delta = __fmul_rn(__fsub_rn(key_comp, sc.keyCompLevel), 0.0078125f);
            sm.recompute_routing = (sc.oldRouting != sc.newRouting);
            if (sm.recompute_routing)
                sc.oldRouting = sc.newRouting;
            sm.draw_bar_change = sc.drawBarChange;
            sm.vib_on = (sc.oldRouting & RT_VIB) != 0;
            sm.perc_on = (sc.oldRouting & RT_PERC) != 0;
            sm.n_ins = 0;
        }
        __syncthreads();
        const bool vib_on = sm.vib_on != 0, perc_on = sm.perc_on != 0;

        /* -- P1: data-independent recurrences on single lanes of warps 1..3, while the
         *        active list is turned into core-program records -- */
        if (tid == 32) {
            gain_ramps(sc, sm, perc_on);
        } else if ((tid == 64 || tid == 65) && whirl_run) {
            rotor_lane(sc, wh, sm, tid == 64); /* horn and drum rotor: two lanes in lockstep */
        }
        active_list_pass(st, tg, sm, tid);
        __syncthreads();

        /* -- P2: additive synthesis, thread = sample; scanner contributions -- */
        float swl, vib, prc;
        core_interpreter(tg, sm, tid, swl, vib, prc);
        sm.vib_in[tid] = vib;
        sm.prc[tid] = prc;
        int vbad = 0;
        if (vib_on)
            vbad = vibrato_contrib(sc, tg, sm, tid, vib);
        vbad = __syncthreads_or(vbad);

        /* -- P3: scanner gather (thread = slot), percussion high-pass, mixdown -- */
        float vib_y = 0.0f;
        if (vib_on) {
            if (!vbad) {
                const float v = vibrato_gather_slot(sc, sm, tid);
                float v2 = 0.0f;
                const bool second = tid + TBF_FRAGMENT < VWINDOW;
                if (second)
                    v2 = vibrato_gather_slot(sc, sm, tid + TBF_FRAGMENT);
                vib_y = sc.vibMixed ? __fmul_rn(__fadd_rn(vib, v), (float)0.7071067811865475) : v;
                __syncthreads(); /* every slot gathered before any ring slot changes */
                sm.vring[(sc.vibOutPos + (uint32_t)tid) & VRING_MASK] = 0.0f;
                if (second)
                    sm.vring[(sc.vibOutPos + (uint32_t)(tid + TBF_FRAGMENT)) & VRING_MASK] = v2;
            } else {
                if (tid == 0)
                    vibrato_serial(sc, tg, sm);
                __syncthreads();
                vib_y = sm.vib_y[tid];
            }
        }
        float prc_hp = 0.0f;
        if (perc_on) /* in-place differencer, src/tonegen.cpp:3719-3732 */
            prc_hp = __fsub_rn(tid ? sm.prc[tid - 1] : sc.pz, prc);
        float mono;
        {
            const float kc = sm.kc[tid];
            if (perc_on) {
                const float og = __fmul_rn(sc.swellPedalGain, sc.percDrawbarGain);
                const float pterm = __fmul_rn(prc_hp, sm.pg[tid]);
                const float acc = vib_on ? __fadd_rn(__fadd_rn(swl, vib_y), pterm) : __fadd_rn(swl, pterm);
                mono = __fmul_rn(__fmul_rn(og, kc), acc);
            } else {
                const float acc = vib_on ? __fadd_rn(swl, vib_y) : swl;
                mono = __fmul_rn(__fmul_rn(sc.swellPedalGain, kc), acc);
            }
        }
        sm.mono[tid] = mono;
        __syncthreads();

        /* -- P4: end-of-fragment tone generator scalars + horn filters A,B (lane 0) -- */
        if (tid == 0) {
            if (vib_on) {
                sc.vibStator = (sc.vibStator + TBF_FRAGMENT * tg->vib_stator_increment) & 0x07FFFFFFu;
                sc.vibOutPos = (sc.vibOutPos + TBF_FRAGMENT) & 0x3FFu;
            }
            if (perc_on) {
                sc.outputGain = __fmul_rn(sc.swellPedalGain, sc.percDrawbarGain);
                sc.pz = sm.prc[TBF_FRAGMENT - 1];
            }
            if (sc.upperKeyCount == 0)
                sc.percEnvGain = sc.percEnvGainReset;
        }
        if (tid < 2 && whirl_run)
            horn_filter_lanes(sc, wh, sm, tid);

        float o_l, o_r;
        if (!whirl_run) {
            if (whirl_off || !whirl3) {
                o_l = o_r = mono;
            } else { /* whirl.bypass=1 seen through whirlProc3's mic mix */
                o_l = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(mono, wh->mic_hll), __fmul_rn(mono, wh->mic_hlr)),
                                          __fmul_rn(0.0f, wh->mic_dll)), __fmul_rn(0.0f, wh->mic_dlr));
                o_r = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(mono, wh->mic_hrl), __fmul_rn(mono, wh->mic_hrr)),
                                          __fmul_rn(0.0f, wh->mic_drl)), __fmul_rn(0.0f, wh->mic_drr));
            }
        } else {
            __syncthreads();
            /* -- P5: the two reflection filters of each path (2-tap FIRs in double);
             *        stage 1 of sample i-1 is recomputed locally to save a barrier -- */
            {
                const float h0 = sm.hx[0][8 + tid], h0m = tid ? sm.hx[0][7 + tid] : sm.zs[0];
                const float h1 = filter_c(h0, h0m);
                const float h1m = tid ? filter_c(h0m, tid > 1 ? sm.hx[0][6 + tid] : sm.zs[0]) : sm.zs[1];
                sm.hx[1][8 + tid] = h1;
                sm.hx[2][8 + tid] = filter_c(h1, h1m);
                const float d0 = sm.dx[0][1 + tid], d0m = tid ? sm.dx[0][tid] : sm.zs[2];
                const float d1 = filter_c(d0, d0m);
                const float d1m = tid ? filter_c(d0m, tid > 1 ? sm.dx[0][tid - 1] : sm.zs[2]) : sm.zs[3];
                sm.dx[1][1 + tid] = d1;
                sm.dx[2][1 + tid] = filter_c(d1, d1m);
            }
            __syncthreads();
            /* -- P6: tap geometry + FIRs, thread = sample -- */
            whirl_taps(wh, sm, tid, outpos0);
            __syncthreads();
            int wbad = 1;
            if (gather_ok)
                wbad = whirl_check_taps(wh, sm, tid, outpos0);
            wbad = __syncthreads_or(wbad);
            /* -- P7: delay lines: ordered parallel scatter or the serial fallback -- */
            if (!wbad) {
                /* slots nearer than the line's shortest delay cannot be written in this
                 * fragment: read and clear them first (this also frees the ring cells
                 * that slots >= ring_size ahead alias with) */
#pragma unroll
                for (int line = 0; line < 4; ++line)
                    if (tid < wh->line_dmin[line]) {
                        float *c = wring + line * ring_size + ((outpos0 + (uint32_t)tid) & rmask);
                        sm.lineout[line][tid] = *c;
                        *c = 0.0f;
                    }
                __syncthreads();
#pragma unroll
                for (int k = 0; k < 3; ++k) {
#pragma unroll
                    for (int line = 0; line < 4; ++line)
                        whirl_tap_pass(sm, wring + line * ring_size, rmask, wh->tap_order[line][k], tid);
                    __syncthreads();
                }
#pragma unroll
                for (int line = 0; line < 4; ++line)
                    if (tid >= wh->line_dmin[line]) {
                        float *c = wring + line * ring_size + ((outpos0 + (uint32_t)tid) & rmask);
                        sm.lineout[line][tid] = *c;
                        *c = 0.0f;
                    }
            } else {
                if (gather_ok && tid == 0)
                    sc.errorFlags |= TBF_IFLAG_GATHER_FALLBACK;
                if ((tid & 31) == 0)
                    scatter_lane(sm, wring + (tid >> 5) * ring_size, rmask, sm.lineout[tid >> 5],
                                 tid >> 5, outpos0);
            }
            __syncthreads();
            /* -- P8: drum filters, one lane per channel in two different warps -- */
            if (tid == 64 || tid == 65) /* two lanes of one warp in lockstep */
                drum_filter_lane(sc, wh, sm, 2 + (tid & 1));
            __syncthreads();
            /* -- P9: output mix -- */
            const float hl = sm.lineout[0][tid], hr = sm.lineout[1][tid];
            const float yl = sm.lineout[2][tid], yr = sm.lineout[3][tid];
            const float leak = sm.leak[tid];
            const float hpl = __fmul_rn(wh->horn_level, hl), hpr = __fmul_rn(wh->horn_level, hr);
            if (!whirl3) {
                o_l = __fadd_rn(__fadd_rn(yl, hpl), leak);
                o_r = __fadd_rn(__fadd_rn(yr, hpr), leak);
            } else {
                const float ohl = __fadd_rn(hpl, leak), ohr = __fadd_rn(hpr, leak);
                o_l = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(ohl, wh->mic_hll), __fmul_rn(ohr, wh->mic_hlr)),
                                          __fmul_rn(yl, wh->mic_dll)), __fmul_rn(yr, wh->mic_dlr));
                o_r = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(ohl, wh->mic_hrl), __fmul_rn(ohr, wh->mic_hrr)),
                                          __fmul_rn(yl, wh->mic_drl)), __fmul_rn(yr, wh->mic_drr));
            }
            /* carry FIR histories / filter states into the next fragment */
            float keep = 0.f;
            if (tid < 24)
                keep = sm.hx[tid >> 3][TBF_FRAGMENT + (tid & 7)];
            float zc = 0.f;
            if (tid >= 32 && tid < 36) {
                const int k = tid - 32;
                zc = (k == 0) ? sm.hx[0][8 + 127] : (k == 1) ? sm.hx[1][8 + 127]
                     : (k == 2) ? sm.dx[0][128] : sm.dx[1][128];
            }
            __syncthreads();
            if (tid < 24)
                sm.hx[tid >> 3][tid & 7] = keep;
            if (tid >= 32 && tid < 36)
                sm.zs[tid - 32] = isnan(zc) ? 0.0f : zc; /* src/whirl.cpp:1627-1630 */
            outpos0 = (outpos0 + TBF_FRAGMENT) & GLOBAL_WBUF_MASK;
        }

        if (store) {
            out_l[(size_t)f * TBF_FRAGMENT + tid] = o_l;
            out_r[(size_t)f * TBF_FRAGMENT + tid] = o_r;
        }
        const size_t o = (size_t)(p.frag_base + f) * TBF_FRAGMENT + tid; /* position in the batch */
        csum += mix64(((unsigned long long)(uint32_t)o << 32) | __float_as_uint(o_l));
        csum += mix64((((unsigned long long)(uint32_t)o << 32) | __float_as_uint(o_r)) ^ 0x5555555555555555ull);
        __syncthreads();
    }

    /* ---- write the staged state back ---- */
    if (tid < 24) {
        const int s = tid >> 3, k = tid & 7;
        sc.adx[s][(sc.adi[s] + k) & 7] = sm.hx[s][7 - k];
    }
    if (tid < 4)
        sc.z[tid] = sm.zs[tid];
    if (tid == 0) {
        sc.outpos = outpos0;
        p.ev_cursor[slot] = ev_cur;
    }
    __syncthreads();
    for (int i = tid; i < (int)(sizeof(InstScalars) / 4); i += TBF_THREADS)
        reinterpret_cast<uint32_t *>(&st->sc)[i] = reinterpret_cast<const uint32_t *>(&sm.sc)[i];
    /* outside the live windows every ring slot is zero (cleared when read) */
    for (int s = tid; s < TBF_VIB_BUF; s += TBF_THREADS) {
        const uint32_t g = (sc.vibOutPos + (uint32_t)s) & 0x3FFu;
        st->vibBuffer[g] = (s < VRING) ? sm.vring[g & VRING_MASK] : 0.0f;
    }
    for (int line = 0; line < 4; ++line)
        for (uint32_t s = tid; s < TBF_WHIRL_BUF; s += TBF_THREADS) {
            const uint32_t g = (outpos0 + s) & GLOBAL_WBUF_MASK;
            st->wbuf[line][g] = (s < ring_size) ? wring[line * ring_size + (g & rmask)] : 0.0f;
        }
    if (p.checksums) {
        atomicAdd(&sm.csum, csum);
        __syncthreads();
        if (tid == 0)
            p.checksums[slot] += sm.csum; /* zeroed by the host; accumulates over chunks */
    }
}

__global__ void tbf_replicate_kernel(InstState *states, uint32_t first, uint32_t count,
                                     const InstState *tmpl)
{
    const uint32_t words = sizeof(InstState) / 4;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(tmpl);
    for (uint32_t inst = blockIdx.x; inst < count; inst += gridDim.x) {
        uint32_t *dst = reinterpret_cast<uint32_t *>(states + first + inst);
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x)
            dst[i] = src[i];
    }
}

} /* namespace */

size_t tbf_render_smem_bytes(int ring_size) { return sizeof(Smem) + (size_t)4 * ring_size * sizeof(float); }

int tbf_launch_render(const RenderParams &p, int ring_size, void *stream)
{
    static int attr_set_for = 0;
    const size_t smem = tbf_render_smem_bytes(ring_size);
    if (attr_set_for < (int)smem) {
        cudaError_t e = cudaFuncSetAttribute(tbf_render_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        attr_set_for = (int)smem;
    }
    tbf_render_kernel<<<p.n_instances, TBF_THREADS, smem, (cudaStream_t)stream>>>(p, (uint32_t)ring_size);
    return (int)cudaGetLastError();
}

int tbf_launch_replicate(InstState *states, uint32_t first, uint32_t count, const InstState *dev_tmpl,
                         void *stream)
{
    if (!count)
        return 0;
    const uint32_t grid = count < 1184 ? count : 1184; /* 148 SMs x 8 */
    tbf_replicate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(states, first, count, dev_tmpl);
    return (int)cudaGetLastError();
}
