/* tbf_kernels.cu -- sm_100a kernels for the tonewheel -> scanner -> Leslie path.
 *
 * One CTA of 128 threads owns one organ instance for a whole batch of
 * fragments: thread n is sample n of the 128-sample fragment for every stage
 * that has no recurrence (additive wheel sum, envelopes, percussion high-pass,
 * Leslie tap geometry + 5-tap FIRs, output mixes), and single lanes in
 * different warps run the short serial recurrences concurrently (biquads,
 * rotor angles, percussion / key-compression ramps, ordered delay-line
 * scatter).  Delay lines live in shared memory for the whole batch.
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
#define WBUF_MASK (TBF_WHIRL_BUF - 1)

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
    float wbuf[4][TBF_WHIRL_BUF];
    float vibbuf[TBF_VIB_BUF];
    CoreRec rec[TBF_NOF_WHEELS + 1];
    uint8_t removed[TBF_NOF_WHEELS + 3];
    uint16_t removed_wheel[TBF_NOF_WHEELS + 1];
    float vib_in[TBF_FRAGMENT], vib_y[TBF_FRAGMENT], prc[TBF_FRAGMENT];
    float pg[TBF_FRAGMENT], kc[TBF_FRAGMENT], mono[TBF_FRAGMENT];
    /* whirl */
    double hang[TBF_FRAGMENT], dang[TBF_FRAGMENT];
    float hx[3][8 + TBF_FRAGMENT]; /* horn stage inputs with 8 samples of history */
    float dx[3][1 + TBF_FRAGMENT]; /* drum stage inputs with 1 sample of history  */
    float leak[TBF_FRAGMENT];
    uint16_t tapn[12][TBF_FRAGMENT];
    float tapa[12][TBF_FRAGMENT], tapb[12][TBF_FRAGMENT];
    float lineout[4][TBF_FRAGMENT]; /* HL HR raw reads, DL DR after the drum biquad */
    unsigned long long csum;
    int n_rec, n_ins, recompute_routing, draw_bar_change, vib_on, perc_on;
    float key_comp_delta;
    float zs[4]; /* FILTER_C states z[0..3], NaN-scrubbed per fragment */
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

__device__ void key_off(InstState *st, const DevTonegenTables *tg, int key)
{
    if (key < 0 || key >= TBF_MAX_KEYS)
        return;
    if (!st->activeKeys[key])
        return;
    st->activeKeys[key] = 0;
    if (key < 128)
        st->upperKeyCount--;
    st->keyDownCount--;
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

__device__ void key_on(InstState *st, const DevTonegenTables *tg, int key)
{
    if (key < 0 || key >= TBF_MAX_KEYS)
        return;
    if (st->activeKeys[key])
        key_off(st, tg, key);
    st->activeKeys[key] = 1;
    if (key < 128)
        st->upperKeyCount++;
    st->keyDownCount++;
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0; e < e1; ++e) {
        const int w = tg->contrib_wheel[e];
        const int b = tg->contrib_bus[e];
        OscState *o = &st->osc[w];
        if (o->refCount == 0) {
            o->rflags = OR_ADD;
            if (o->aclPos == -1) {
                o->aclPos = st->activeOscLEnd;
                st->activeOscList[st->activeOscLEnd++] = w;
            }
        } else {
            o->rflags |= ORF_MODIFIED;
        }
        float *bl = &st->busLevel[w * TBF_NOF_BUSES + b];
        *bl = *bl + tg->contrib_level[e];
        o->refCount += 1;
    }
}

__device__ void perc_resets(InstState *st, const DevTonegenTables *tg)
{
    if (st->percIsFast)
        st->percEnvGainDecay = st->percIsSoft ? tg->perc_decay_fast_soft : tg->perc_decay_fast_norm;
    else
        st->percEnvGainDecay = st->percIsSoft ? tg->perc_decay_slow_soft : tg->perc_decay_slow_norm;
}

__device__ void use_rev_option(InstState *st, const DevWhirlTables *wh, int n, int signals)
{
    const int i = n % 9;
    st->hornTarget = wh->rev_horn_target[i];
    st->drumTarget = wh->rev_drum_target[i];
    if (st->hornIncr < st->hornTarget)
        st->hornAcDc = 1;
    else if (st->hornTarget < st->hornIncr)
        st->hornAcDc = -1;
    if (st->drumIncr < st->drumTarget)
        st->drumAcDc = 1;
    else if (st->drumTarget < st->drumIncr)
        st->drumAcDc = -1;
    if (signals & 2) {
        const int hr = (n / 3) % 3;
        st->revSelect = (hr == 2) ? 2 : (hr == 1) ? 0 : 1; /* WHIRL_FAST / SLOW / STOP */
    }
}

__device__ void apply_event(InstState *st, const DevTonegenTables *tg, const DevWhirlTables *wh,
                            const tbf_event &ev)
{
    const int a = ev.a, b = ev.b;
    switch (ev.type) {
    case TBF_EV_KEY_ON:
        key_on(st, tg, a);
        break;
    case TBF_EV_KEY_OFF:
        key_off(st, tg, a);
        break;
    case TBF_EV_DRAWBAR:
        if (a < 0 || a >= TBF_NOF_BUSES || b < 0 || b >= 9) { /* reference asserts */
            st->errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        st->drawBarChange = 1;
        if (a == tg->perc_trigger_bus) {
            st->percTrigRestore = b;
            if (st->percEnabled)
                break;
        }
        st->drawBarGain[a] = tg->draw_bar_level[a][b];
        break;
    case TBF_EV_PERC_ENABLE:
        if (a) {
            st->newRouting |= RT_PERC;
            if (-1 < tg->perc_trigger_bus) {
                st->drawBarGain[tg->perc_trigger_bus] = 0.0f;
                st->drawBarChange = 1;
            }
        } else {
            st->newRouting &= ~RT_PERC;
            if (-1 < tg->perc_trigger_bus) {
                st->drawBarGain[tg->perc_trigger_bus] =
                    tg->draw_bar_level[tg->perc_trigger_bus][st->percTrigRestore];
                st->drawBarChange = 1;
            }
        }
        st->percEnabled = a;
        break;
    case TBF_EV_PERC_SOFT:
        st->percIsSoft = a;
        st->percEnvGainReset = __fmul_rn(
            tg->perc_env_scaling, a ? tg->perc_env_gain_reset_soft : tg->perc_env_gain_reset_norm);
        st->percDrawbarGain = a ? tg->perc_drawbar_soft_gain : tg->perc_drawbar_normal_gain;
        perc_resets(st, tg);
        break;
    case TBF_EV_PERC_FAST:
        st->percIsFast = a;
        perc_resets(st, tg);
        break;
    case TBF_EV_PERC_FIRST:
        st->percSendBus = a ? tg->perc_send_bus_a : tg->perc_send_bus_b;
        break;
    case TBF_EV_VIB_UPPER:
        if (a)
            st->newRouting |= RT_UPPRVIB;
        else
            st->newRouting &= ~RT_UPPRVIB;
        break;
    case TBF_EV_VIB_LOWER:
        if (a)
            st->newRouting |= RT_LOWRVIB;
        else
            st->newRouting &= ~RT_LOWRVIB;
        break;
    case TBF_EV_VIB_KNOB:
        if (a >= 0 && a <= 5) { /* V1 C1 V2 C2 V3 C3 */
            st->vibTable = a >> 1;
            st->vibMixed = (a & 1) ? 0x80 : 0;
        }
        break;
    case TBF_EV_REV_OPTION:
        if (a < 0) {
            st->errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        use_rev_option(st, wh, a, 2);
        break;
    case TBF_EV_REV_SELECT:
        if (a < 0) {
            st->errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        st->revSelect = a % 3;
        use_rev_option(st, wh, wh->revselects[st->revSelect], 1);
        break;
    case TBF_EV_SWELL:
        st->swellPedalGain =
            (float)(((double)tg->output_level_trim * (double)(unsigned char)a) / 127.0);
        break;
    default:
        st->errorFlags |= TBF_IFLAG_BAD_EVENT;
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
    const int n_act = st->activeOscLEnd;
    const uint32_t routing = st->oldRouting; /* already updated by thread 0 */
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
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], st->drawBarGain[d]));
                o->sumUpper = sum;
                sum = 0.0f;
                for (int d = 9; d < 18; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], st->drawBarGain[d]));
                o->sumLower = sum;
                sum = 0.0f;
                for (int d = 18; d < 27; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], st->drawBarGain[d]));
                o->sumPedal = sum;
                reroute = true;
            }
            if (reroute || recompute_routing) {
                float percn = 0.0f;
                if (routing & RT_PERC)
                    percn = st->busLevel[w * TBF_NOF_BUSES + st->percSendBus];
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
        st->statFragments += 1;
        st->statActiveWheels += (unsigned long long)n_act;
        st->drawBarChange = 0;
        if (sm.n_ins > 256)
            st->errorFlags |= TBF_IFLAG_COREPGM_OVERFLOW;
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
            st->activeOscLEnd = l_end;
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
/* scanner vibrato, ordered scatter by one lane (src/vibrato.cpp:365-411) */
/* ------------------------------------------------------------------ */
__device__ void vibrato_serial(InstState *st, const DevTonegenTables *tg, Smem &sm)
{
    const float fnorm = (float)(1.0 / 65536.0);
    const float mixnorm = (float)0.7071067811865475;
    const uint32_t *tab = tg->vib_offset_table[st->vibTable];
    const uint32_t inc = tg->vib_stator_increment;
    const int mixed = st->vibMixed;
    uint32_t stator = st->vibStator, out_pos = st->vibOutPos;
    float *buf = sm.vibbuf;
    for (int i = 0; i < TBF_FRAGMENT; ++i) {
        const float x = sm.vib_in[i];
        const uint32_t j = ((out_pos << 16) + __ldg(&tab[stator >> 16])) & 0x03FFFFFFu;
        const uint32_t h = j >> 16;
        const uint32_t k = (h + 1) & 0x3FFu;
        const float f = __fmul_rn(fnorm, (float)(j & 0xFFFFu));
        const float g = __fmul_rn(f, x);
        buf[h] = __fadd_rn(buf[h], __fsub_rn(x, g));
        buf[k] = __fadd_rn(buf[k], g);
        float y;
        if (mixed)
            y = __fmul_rn(__fadd_rn(x, buf[out_pos]), mixnorm);
        else
            y = buf[out_pos];
        sm.vib_y[i] = y;
        buf[out_pos] = 0.0f;
        out_pos = (out_pos + 1) & 0x3FFu;
        stator = (stator + inc) & 0x07FFFFFFu;
    }
    st->vibStator = stator;
    st->vibOutPos = out_pos;
}

/* percussion-gain decay and key-compression chase: two scalar recurrences
 * carried across fragments (src/tonegen.cpp:3733-3771), one lane */
__device__ void gain_ramps(InstState *st, Smem &sm, bool perc_on)
{
    float kc = st->keyCompLevel;
    const float delta = sm.key_comp_delta;
    if (perc_on) {
        float pg = st->percEnvGain;
        const float decay = st->percEnvGainDecay;
        for (int i = 0; i < TBF_FRAGMENT; ++i) {
            sm.pg[i] = pg;
            sm.kc[i] = kc;
            pg = __fmul_rn(pg, decay);
            kc = __fadd_rn(kc, delta);
        }
        st->percEnvGain = pg;
    } else {
        for (int i = 0; i < TBF_FRAGMENT; ++i) {
            sm.kc[i] = kc;
            kc = __fadd_rn(kc, delta);
        }
    }
    st->keyCompLevel = kc;
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
    for (int i = 0; i < TBF_FRAGMENT; ++i) {
        out[i] = a;
        a = angle_step(a, incr);
    }
    *final_angle = a;
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

/* horn input filters A and B in series, one lane (src/whirl.cpp:1513-1525) */
__device__ void horn_filter_lane(InstState *st, const DevWhirlTables *wh, Smem &sm)
{
    float c0[5], c1[5];
    for (int k = 0; k < 5; ++k) {
        c0[k] = wh->iir[0][k];
        c1[k] = wh->iir[1][k];
    }
    float a0 = st->iirz[0][0], a1 = st->iirz[0][1];
    float b0 = st->iirz[1][0], b1 = st->iirz[1][1];
    const float leakage = wh->leakage;
    for (int i = 0; i < TBF_FRAGMENT; ++i) {
        const float x = (float)__dadd_rn((double)sm.mono[i], 1e-14); /* DENORMAL_HACK */
        sm.dx[0][1 + i] = x;
        float y = eq_iir(c0, a0, a1, x);
        y = eq_iir(c1, b0, b1, y);
        sm.hx[0][8 + i] = y;
        sm.leak[i] = __fmul_rn(y, leakage);
    }
    /* NaN scrub after the loop (EQ_IIR_NAN, src/whirl.cpp:1622-1625) */
    st->iirz[0][0] = isnan(a0) ? 0.0f : a0;
    st->iirz[0][1] = isnan(a1) ? 0.0f : a1;
    st->iirz[1][0] = isnan(b0) ? 0.0f : b0;
    st->iirz[1][1] = isnan(b1) ? 0.0f : b1;
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
/* of every tap to shared memory; the ordered accumulation into the     */
/* delay lines is done by scatter_lane().                               */
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
    slot = f2u_x86(r) & WBUF_MASK;
}

__device__ void whirl_taps(const DevWhirlTables *wh, Smem &sm, int i, uint32_t outpos0)
{
    const uint32_t outpos = (outpos0 + (uint32_t)i) & WBUF_MASK;
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

/* One lane per delay line (0 HL, 1 HR, 2 DL, 3 DR): per sample, accumulate the
 * line's three taps in source order, then read and clear the output slot; the
 * drum lines run their biquad on the value read (src/whirl.cpp:1540-1616). */
__device__ void scatter_lane(InstState *st, const DevWhirlTables *wh, Smem &sm, int line,
                             uint32_t outpos0)
{
    float *buf = sm.wbuf[line];
    const int tap0 = (line < 2) ? (line & 1) : 6 + (line & 1);
    float c[5];
    float z0 = 0.f, z1 = 0.f;
    if (line >= 2) {
        for (int k = 0; k < 5; ++k)
            c[k] = wh->iir[line][k];
        z0 = st->iirz[line][0];
        z1 = st->iirz[line][1];
    }
    for (int i = 0; i < TBF_FRAGMENT; ++i) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const int p = tap0 + 2 * s;
            const uint32_t n = sm.tapn[p][i];
            buf[n] = __fadd_rn(buf[n], sm.tapa[p][i]);
            const uint32_t n1 = (n + 1) & WBUF_MASK;
            buf[n1] = __fadd_rn(buf[n1], sm.tapb[p][i]);
        }
        const uint32_t op = (outpos0 + (uint32_t)i) & WBUF_MASK;
        float v = buf[op];
        buf[op] = 0.0f;
        if (line >= 2)
            v = eq_iir(c, z0, z1, v);
        sm.lineout[line][i] = v;
    }
    if (line >= 2) {
        st->iirz[line][0] = isnan(z0) ? 0.0f : z0;
        st->iirz[line][1] = isnan(z1) ? 0.0f : z1;
    }
}

/* ------------------------------------------------------------------ */
/* the batched render kernel: grid = instances, block = 128 threads     */
/* ------------------------------------------------------------------ */
__global__ void __launch_bounds__(TBF_THREADS)
tbf_render_kernel(const RenderParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
    const int tid = threadIdx.x;
    const uint32_t slot = blockIdx.x;
    const uint32_t id = p.ids ? p.ids[slot] : slot;
    InstState *st = p.states + id;
    const DevTonegenTables *tg = p.tg + st->tgTables;
    const DevWhirlTables *wh = p.wh + st->whTables;
    const bool whirl_off = (p.flags & TBF_RENDER_WHIRL_BYPASS) != 0;
    const bool whirl3 = (p.flags & TBF_RENDER_WHIRLPROC3) != 0;
    const bool whirl_run = !whirl_off && !wh->bypass;
    const bool store = !(p.flags & TBF_RENDER_NO_OUTPUT);
    float *out_l = p.out_l + (size_t)slot * p.out_stride;
    float *out_r = p.out_r + (size_t)slot * p.out_stride;

    /* ---- stage the delay lines and FIR histories in shared memory ---- */
    for (int i = tid; i < TBF_VIB_BUF; i += TBF_THREADS)
        sm.vibbuf[i] = st->vibBuffer[i];
    for (int i = tid; i < 4 * TBF_WHIRL_BUF; i += TBF_THREADS)
        (&sm.wbuf[0][0])[i] = (&st->wbuf[0][0])[i];
    if (tid < 24) {
        const int s = tid >> 3, k = tid & 7;
        sm.hx[s][7 - k] = st->adx[s][(st->adi[s] + k) & 7];
    }
    if (tid < 4)
        sm.zs[tid] = st->z[tid];
    if (tid == 0)
        sm.csum = 0ull;
    uint32_t outpos0 = st->outpos;
    uint64_t ev_cur = p.ev_cursor[slot];
    const uint64_t ev_end = p.ev_start[slot + 1];
    unsigned long long csum = 0ull;
    __syncthreads();

    for (uint32_t f = 0; f < p.n_fragments; ++f) {
        /* -- S0: control events that precede this fragment, routing, key compression -- */
        if (tid == 0) {
            while (ev_cur < ev_end && p.events[ev_cur].fragment <= p.frag_base + f) {
                apply_event(st, tg, wh, p.events[ev_cur]);
                ++ev_cur;
            }
            int kdc = st->keyDownCount;
            if (kdc > 127 || kdc < 0) { /* reference reads past keyCompTable[128] */
                st->errorFlags |= TBF_IFLAG_KEYCOMP_OOB;
                kdc = kdc < 0 ? 0 : 127;
            }
            const float key_comp = tg->key_comp_table[kdc];
            /* (keyComp - level) / 128.f: dividing by 2^7 and multiplying by 2^-7 round identically */
            sm.key_comp_delta = __fmul_rn(__fsub_rn(key_comp, st->keyCompLevel), 0.0078125f);
            sm.recompute_routing = (st->oldRouting != st->newRouting);
            if (sm.recompute_routing)
                st->oldRouting = st->newRouting;
            sm.draw_bar_change = st->drawBarChange;
            sm.vib_on = (st->oldRouting & RT_VIB) != 0;
            sm.perc_on = (st->oldRouting & RT_PERC) != 0;
            sm.n_ins = 0;
        }
        __syncthreads();

        /* -- S1: active list -> core program records -- */
        active_list_pass(st, tg, sm, tid);
        __syncthreads();
        const bool vib_on = sm.vib_on != 0, perc_on = sm.perc_on != 0;

        /* -- S2: additive synthesis, thread = sample -- */
        float swl, vib, prc;
        core_interpreter(tg, sm, tid, swl, vib, prc);
        sm.vib_in[tid] = vib;
        sm.prc[tid] = prc;
        __syncthreads();

        /* -- S3: serial lanes in four different warps + percussion high-pass -- */
        if (tid == 0) {
            if (vib_on)
                vibrato_serial(st, tg, sm);
        } else if (tid == 32) {
            gain_ramps(st, sm, perc_on);
        } else if (tid == 64 && whirl_run) {
            const RotorConst rc = {10.f, 60.f, 1};
            int acdc = st->hornAcDc;
            double incr = st->hornIncr, ang = st->hornAngle;
            const int brake = rotor_ramp(wh, rc, &acdc, &incr, &ang, st->hornTarget,
                                         wh->hn_brake_pos, wh->one_minus_l[0], wh->one_minus_l[1]);
            double fin;
            angle_lane(ang, incr, sm.hang, &fin);
            st->hornAcDc = acdc;
            st->hornIncr = brake ? 0.0 : incr;
            st->hornAngle = fin;
        } else if (tid == 96 && whirl_run) {
            const RotorConst rc = {8.f, 100.f, 0};
            int acdc = st->drumAcDc;
            double incr = st->drumIncr, ang = st->drumAngle;
            const int brake = rotor_ramp(wh, rc, &acdc, &incr, &ang, st->drumTarget,
                                         wh->dr_brake_pos, wh->one_minus_l[2], wh->one_minus_l[3]);
            double fin;
            angle_lane(ang, incr, sm.dang, &fin);
            st->drumAcDc = acdc;
            st->drumIncr = brake ? 0.0 : incr;
            st->drumAngle = fin;
        }
        float prc_hp = 0.0f;
        if (perc_on) /* in-place differencer, src/tonegen.cpp:3719-3732 */
            prc_hp = __fsub_rn(tid ? sm.prc[tid - 1] : st->pz, prc);
        __syncthreads();

        /* -- S4: mixdown (src/tonegen.cpp:3733-3777) -- */
        float mono;
        {
            const float kc = sm.kc[tid];
            if (perc_on) {
                const float og = __fmul_rn(st->swellPedalGain, st->percDrawbarGain);
                const float pterm = __fmul_rn(prc_hp, sm.pg[tid]);
                const float acc = vib_on ? __fadd_rn(__fadd_rn(swl, sm.vib_y[tid]), pterm)
                                         : __fadd_rn(swl, pterm);
                mono = __fmul_rn(__fmul_rn(og, kc), acc);
            } else {
                const float acc = vib_on ? __fadd_rn(swl, sm.vib_y[tid]) : swl;
                mono = __fmul_rn(__fmul_rn(st->swellPedalGain, kc), acc);
            }
        }
        sm.mono[tid] = mono;
        __syncthreads();
        if (tid == 0) {
            if (perc_on) {
                st->outputGain = __fmul_rn(st->swellPedalGain, st->percDrawbarGain);
                st->pz = sm.prc[TBF_FRAGMENT - 1];
            }
            if (st->upperKeyCount == 0)
                st->percEnvGain = st->percEnvGainReset;
        }

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
            /* -- S5: horn filters A,B: serial lane -- */
            if (tid == 0)
                horn_filter_lane(st, wh, sm);
            __syncthreads();
            /* -- S6: reflection filters (2-tap FIRs in double) -- */
            sm.hx[1][8 + tid] = filter_c(sm.hx[0][8 + tid], tid ? sm.hx[0][7 + tid] : sm.zs[0]);
            sm.dx[1][1 + tid] = filter_c(sm.dx[0][1 + tid], tid ? sm.dx[0][tid] : sm.zs[2]);
            __syncthreads();
            sm.hx[2][8 + tid] = filter_c(sm.hx[1][8 + tid], tid ? sm.hx[1][7 + tid] : sm.zs[1]);
            sm.dx[2][1 + tid] = filter_c(sm.dx[1][1 + tid], tid ? sm.dx[1][tid] : sm.zs[3]);
            __syncthreads();
            /* -- S7: tap geometry + FIRs, thread = sample -- */
            whirl_taps(wh, sm, tid, outpos0);
            __syncthreads();
            /* -- S8: ordered scatter / read / drum biquads: one lane per line -- */
            if ((tid & 31) == 0)
                scatter_lane(st, wh, sm, tid >> 5, outpos0);
            __syncthreads();
            /* -- S9: output mix -- */
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
            outpos0 = (outpos0 + TBF_FRAGMENT) & WBUF_MASK;
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
    for (int i = tid; i < TBF_VIB_BUF; i += TBF_THREADS)
        st->vibBuffer[i] = sm.vibbuf[i];
    for (int i = tid; i < 4 * TBF_WHIRL_BUF; i += TBF_THREADS)
        (&st->wbuf[0][0])[i] = (&sm.wbuf[0][0])[i];
    if (tid < 24) {
        const int s = tid >> 3, k = tid & 7;
        st->adx[s][(st->adi[s] + k) & 7] = sm.hx[s][7 - k];
    }
    if (tid < 4)
        st->z[tid] = sm.zs[tid];
    if (tid == 0) {
        st->outpos = outpos0;
        p.ev_cursor[slot] = ev_cur;
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

size_t tbf_render_smem_bytes() { return sizeof(Smem); }

int tbf_launch_render(const RenderParams &p, void *stream)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(tbf_render_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)sizeof(Smem));
        if (e != cudaSuccess)
            return (int)e;
        attr_set = true;
    }
    tbf_render_kernel<<<p.n_instances, TBF_THREADS, sizeof(Smem), (cudaStream_t)stream>>>(p);
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
