/* ORACLE (test infrastructure, not product code).
 *
 * C-ABI harness around the UNMODIFIED reference engine, compiled in place from
 * /root/reference/src by oracle/Makefile into oracle/_ref/libtbf_ref.so.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs load it.
 *
 * It follows the allocation / init order of the reference's LV2 host
 * (b_synth/lv2.cpp:336-353 allocSynth, :164-193 initSynth) and the per-fragment
 * chain of its hosts (src/clap.cpp:254-258) with overdrive clean (a memcpy,
 * src/overdrive.cpp:329-335) and reverb skipped (BASELINE config 1).
 * Build flags are strict IEEE (no -ffast-math, -ffp-contract=off): SURVEY 6.2.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>

#include "cfgParser.h"
#include "global_inst.h"
#include "midi.h"
#include "overdrive.h"
#include "pgmParser.h"
#include "program.h"
#include "reverb.h"
#include "state.h"
#include "tonegen.h"
#include "vibrato.h"
#include "whirl.h"

#include "tbf_gpu.h"

extern "C" void ref_mts_set_table(const double *freqs128);

/* symbols every reference host has to provide (b_synth/lv2.cpp:135-162) */
const ConfigDoc *mainDoc() { return NULL; }
int mainConfig(ConfigContext *cfg)
{
    (void)cfg;
    return 0;
}
double SampleRateD = 48000.0;

struct ref_inst {
    b_instance inst;
    double rate;
    float bufA[128], tmpL[128], tmpR[128];
    /* workload statistics for bench.py: fragments rendered and the sum over them of the
     * active-list length (the larger of its length before and after the fragment, i.e. with
     * the wheels a key event adds or removes in that fragment still counted) */
    unsigned long long stat_fragments, stat_active_wheels;
};

extern "C" {

void *ref_create(double rate, const double *freqs128, const double *target_ratio9, unsigned seed,
                 const char *cfg_text)
{
    ref_inst *r = (ref_inst *)calloc(1, sizeof(ref_inst));
    if (!r)
        return NULL;
    r->rate = rate;
    SampleRateD = rate;
    if (freqs128)
        ref_mts_set_table(freqs128);

    b_instance *inst = &r->inst;
    inst->state = allocRunningConfig();
    inst->progs = allocProgs();
    inst->reverb = allocReverb();
    inst->whirl = allocWhirl();
    inst->midicfg = allocMidiCfg(inst->state);
    inst->synth = allocTonegen();
    inst->preamp = allocPreamp();
    initControllerTable(inst->midicfg);
    midiPrimeControllerMapping(inst->midicfg);

    if (cfg_text && *cfg_text) {
        char *copy = strdup(cfg_text);
        char *save = NULL;
        int ln = 0;
        for (char *line = strtok_r(copy, "\n", &save); line; line = strtok_r(NULL, "\n", &save)) {
            char buf[2048];
            strncpy(buf, line, sizeof(buf) - 1);
            buf[sizeof(buf) - 1] = 0;
            parseConfigurationLine(inst, "<oracle>", ++ln, buf);
        }
        free(copy);
    }

    /* hosts seed with time(NULL) (src/main.cpp:805, b_synth/lv2.cpp:949); the
     * oracle pins the glibc rand() stream right before the tone generator
     * consumes it (dither src/tonegen.cpp:1450, click envelopes :2577-2660). */
    srand(seed);
    double tr[9];
    if (target_ratio9)
        memcpy(tr, target_ratio9, sizeof(tr));
    initToneGenerator(inst->synth, inst->midicfg, rate, target_ratio9 ? tr : NULL);
    initVibrato(inst->synth, inst->midicfg);
    initPreamp(inst->preamp, inst->midicfg, rate);
    initReverb(inst->reverb, inst->midicfg, rate);
    initWhirl(inst->whirl, inst->midicfg, rate);
    initRunningConfig(inst->state, inst->midicfg);
    initMidiTables(inst->midicfg);
    return r;
}

void ref_destroy(void *h)
{
    ref_inst *r = (ref_inst *)h;
    if (!r)
        return;
    freeReverb(r->inst.reverb);
    freeWhirl(r->inst.whirl);
    freeToneGenerator(r->inst.synth);
    freeMidiCfg(r->inst.midicfg);
    freePreamp(r->inst.preamp);
    freeProgs(r->inst.progs);
    freeRunningConfig(r->inst.state);
    free(r);
}

/* one control event, same numbering as tbf_event_type (include/tbf_gpu.h) */
int ref_event(void *h, int type, int a, int b)
{
    ref_inst *r = (ref_inst *)h;
    struct b_tonegen *t = r->inst.synth;
    struct b_whirl *w = r->inst.whirl;
    switch (type) {
    case TBF_EV_KEY_ON:
        oscKeyOn(t, (short)a, (short)a);
        break;
    case TBF_EV_KEY_OFF:
        oscKeyOff(t, (short)a, (short)a);
        break;
    case TBF_EV_DRAWBAR:
        setDrawBar(t, a, (unsigned int)b);
        break;
    case TBF_EV_PERC_ENABLE:
        setPercussionEnabled(t, a);
        break;
    case TBF_EV_PERC_SOFT:
        setPercussionVolume(t, a);
        break;
    case TBF_EV_PERC_FAST:
        setPercussionFast(t, a);
        break;
    case TBF_EV_PERC_FIRST:
        setPercussionFirst(t, a);
        break;
    case TBF_EV_VIB_UPPER:
        setVibratoUpper(t, a);
        break;
    case TBF_EV_VIB_LOWER:
        setVibratoLower(t, a);
        break;
    case TBF_EV_VIB_KNOB:
        setVibratoFromInt(t, a);
        break;
    case TBF_EV_REV_OPTION:
        useRevOption(w, a, 2);
        break;
    case TBF_EV_REV_SELECT:
        setRevSelect(w, a);
        break;
    case TBF_EV_SWELL:
        /* setSwellPedal1FromMIDI is static in the reference (src/tonegen.cpp:2885-2890);
         * this is its one assignment. */
        t->swellPedalGain = (t->outputLevelTrim * ((double)(unsigned char)a)) / 127.0;
        break;
    /* the controller callbacks below are static in the reference (src/whirl.cpp:883-917);
     * each case is that callback's one assignment */
    case TBF_EV_WHIRL_BRAKE:
        if (a == 0)
            w->hnBrakePos = (double)(unsigned char)b / 127.0;
        else
            w->drBrakePos = (double)(unsigned char)b / 127.0;
        break;
    case TBF_EV_WHIRL_RAMP:
        if (a == 0)
            w->hornAcc = .01 + (double)(unsigned char)b / 80.0;
        else if (a == 1)
            w->hornDec = .01 + (double)(unsigned char)b / 80.0;
        else if (a == 2)
            w->drumAcc = .01 + (double)(unsigned char)b / 14.0;
        else
            w->drumDec = .01 + (double)(unsigned char)b / 14.0;
        break;
    case TBF_EV_HORN_FILTER: {
        /* what setIIRFilter() stores (src/whirl.cpp:167-171), one coefficient at a time */
        static const int idx[5] = {a1, a2, b0, b1, b2};
        float v;
        memcpy(&v, &b, sizeof v);
        ((a >> 3) ? w->hbfw : w->hafw)[idx[a & 7]] = v;
        break;
    }
    default:
        return -1;
    }
    return 0;
}

/* Render n_fragments x 128 samples.  flags = TBF_RENDER_* of tbf_gpu.h.
 * events (may be NULL) must be sorted by fragment; instance field ignored.  */
int ref_render(void *h, unsigned n_fragments, unsigned flags, const tbf_event *events,
               unsigned long long n_events, float *outL, float *outR)
{
    ref_inst *r = (ref_inst *)h;
    unsigned long long e = 0;
    for (unsigned f = 0; f < n_fragments; ++f) {
        while (e < n_events && events[e].fragment == f) {
            ref_event(h, events[e].type, events[e].a, events[e].b);
            ++e;
        }
        float *L = outL + (size_t)f * 128;
        float *R = outR + (size_t)f * 128;
        const int act_before = r->inst.synth->activeOscLEnd;
        oscGenerateFragment(r->inst.synth, r->bufA, 128);
        const int act_after = r->inst.synth->activeOscLEnd;
        r->stat_fragments += 1;
        r->stat_active_wheels += (unsigned long long)(act_before > act_after ? act_before : act_after);
        if (flags & TBF_RENDER_WHIRL_BYPASS) {
            memcpy(L, r->bufA, sizeof(float) * 128);
            memcpy(R, r->bufA, sizeof(float) * 128);
        } else if (flags & TBF_RENDER_WHIRLPROC3) {
            whirlProc3(r->inst.whirl, r->bufA, L, R, r->tmpL, r->tmpR, 128);
        } else {
            whirlProc(r->inst.whirl, r->bufA, L, R, 128);
        }
    }
    return 0;
}

/* The Leslie alone on the caller's mono input (whirlProc / whirlProc3, src/whirl.cpp:1640,
 * :1653): only rotor events are applied.  Oracle for TBF_RENDER_WHIRL_INPUT. */
int ref_whirl_only(void *h, unsigned n_fragments, unsigned flags, const float *in,
                   const tbf_event *events, unsigned long long n_events, float *outL, float *outR)
{
    ref_inst *r = (ref_inst *)h;
    unsigned long long e = 0;
    for (unsigned f = 0; f < n_fragments; ++f) {
        while (e < n_events && events[e].fragment == f) {
            if (events[e].type == TBF_EV_REV_OPTION || events[e].type == TBF_EV_REV_SELECT ||
                events[e].type == TBF_EV_WHIRL_BRAKE || events[e].type == TBF_EV_WHIRL_RAMP ||
                events[e].type == TBF_EV_HORN_FILTER)
                ref_event(h, events[e].type, events[e].a, events[e].b);
            ++e;
        }
        const float *x = in + (size_t)f * 128;
        float *L = outL + (size_t)f * 128;
        float *R = outR + (size_t)f * 128;
        if (flags & TBF_RENDER_WHIRLPROC3)
            whirlProc3(r->inst.whirl, x, L, R, r->tmpL, r->tmpR, 128);
        else
            whirlProc(r->inst.whirl, x, L, R, 128);
    }
    return 0;
}

/* whirlProc2 on the caller's mono input with its separate horn / drum outputs
 * (src/whirl.cpp:1191, src/whirl.h:232-238).  Oracle for TBF_RENDER_STEMS. */
int ref_whirl2(void *h, unsigned n_fragments, const float *in, float *outL, float *outR, float *outHL,
               float *outHR, float *outDL, float *outDR)
{
    ref_inst *r = (ref_inst *)h;
    for (unsigned f = 0; f < n_fragments; ++f) {
        const size_t o = (size_t)f * 128;
        whirlProc2(r->inst.whirl, in + o, outL + o, outR + o, outHL + o, outHR + o, outDL + o, outDR + o,
                   128);
    }
    return 0;
}

/* Render with RAW MIDI input, the way the hosts do it: every message goes through the
 * reference's own parse_raw_midi_data() (src/midi.cpp:1261) right before the fragment it is
 * stamped with (b_synth/lv2.cpp:1130-1136).  msgs = n_msgs x {fragment, size, b0, b1, b2} as
 * uint32, sorted by fragment.  Oracle for the MIDI -> event compiler (include/tbf_midi.h). */
int ref_render_midi(void *h, unsigned n_fragments, unsigned flags, const unsigned *msgs,
                    unsigned long long n_msgs, float *outL, float *outR)
{
    ref_inst *r = (ref_inst *)h;
    unsigned long long e = 0;
    for (unsigned f = 0; f < n_fragments; ++f) {
        while (e < n_msgs && msgs[5 * e] == f) {
            uint8_t b[3] = {(uint8_t)msgs[5 * e + 2], (uint8_t)msgs[5 * e + 3], (uint8_t)msgs[5 * e + 4]};
            parse_raw_midi_data(&r->inst, b, msgs[5 * e + 1]);
            ++e;
        }
        float *L = outL + (size_t)f * 128;
        float *R = outR + (size_t)f * 128;
        oscGenerateFragment(r->inst.synth, r->bufA, 128);
        if (flags & TBF_RENDER_WHIRL_BYPASS) {
            memcpy(L, r->bufA, sizeof(float) * 128);
            memcpy(R, r->bufA, sizeof(float) * 128);
        } else if (flags & TBF_RENDER_WHIRLPROC3) {
            whirlProc3(r->inst.whirl, r->bufA, L, R, r->tmpL, r->tmpR, 128);
        } else {
            whirlProc(r->inst.whirl, r->bufA, L, R, 128);
        }
    }
    return 0;
}

/* Load programme definitions (.pgm syntax) into the instance's programme table with the
 * reference's own parser (loadProgrammeFile, src/pgmParser.cpp:376; the string variant is
 * LV2-only).  0 = parsed. */
int ref_load_programmes(void *h, const char *text)
{
    ref_inst *r = (ref_inst *)h;
    char path[] = "/tmp/tbf_oracle_pgm_XXXXXX";
    const int fd = mkstemp(path);
    if (fd < 0)
        return -1;
    FILE *fp = fdopen(fd, "w");
    fputs(text, fp);
    fclose(fp);
    const int rc = loadProgrammeFile(r->inst.progs, path);
    remove(path);
    return rc;
}

/* the reference's note -> key table of a channel (-1: none), for the key-mapping tests */
int ref_midi_key(void *h, int channel, int note)
{
    ref_inst *r = (ref_inst *)h;
    /* b_midicfg is private to midi.cpp: observe the mapping through the engine instead */
    struct b_tonegen *t = r->inst.synth;
    uint8_t on[3] = {(uint8_t)(0x90 | (channel & 15)), (uint8_t)(note & 127), 100};
    uint8_t off[3] = {(uint8_t)(0x80 | (channel & 15)), (uint8_t)(note & 127), 0};
    int key = -1;
    parse_raw_midi_data(&r->inst, on, 3);
    for (int k = 0; k < MAX_KEYS; ++k)
        if (t->activeKeys[k]) {
            key = k;
            break;
        }
    parse_raw_midi_data(&r->inst, off, 3);
    return key;
}

int ref_active_wheels(void *h) { return ((ref_inst *)h)->inst.synth->activeOscLEnd; }

/* ---- flatten the reference structs into the boundary's table format ---- */

int ref_dump_tonegen_tables(void *h, tbf_tonegen_tables *o)
{
    ref_inst *r = (ref_inst *)h;
    struct b_tonegen *t = r->inst.synth;
    memset(o, 0, sizeof(*o));

    uint64_t total = 0;
    for (int w = 1; w <= NOF_WHEELS; ++w)
        total += t->oscillators[w].lengthSamples;
    float *wd = (float *)malloc(sizeof(float) * (total ? total : 1));
    uint64_t off = 0;
    for (int w = 1; w <= NOF_WHEELS; ++w) {
        size_t n = t->oscillators[w].lengthSamples;
        o->wave_offset[w] = (uint32_t)off;
        o->wave_length[w] = (uint32_t)n;
        if (n)
            memcpy(wd + off, t->oscillators[w].wave, n * sizeof(float));
        off += n;
    }
    o->wave_data = wd;
    o->wave_total = total;

    uint64_t nc = 0;
    for (int k = 0; k < MAX_KEYS; ++k)
        for (ListElement *l = t->keyContrib[k]; l; l = l->next)
            ++nc;
    uint16_t *cw = (uint16_t *)malloc(sizeof(uint16_t) * (nc ? nc : 1));
    uint16_t *cb = (uint16_t *)malloc(sizeof(uint16_t) * (nc ? nc : 1));
    float *cl = (float *)malloc(sizeof(float) * (nc ? nc : 1));
    uint64_t c = 0;
    for (int k = 0; k < MAX_KEYS; ++k) {
        o->contrib_start[k] = (uint32_t)c;
        for (ListElement *l = t->keyContrib[k]; l; l = l->next) {
            cw[c] = (uint16_t)l->u.ssf.sa;
            cb[c] = (uint16_t)l->u.ssf.sb;
            cl[c] = l->u.ssf.fc;
            ++c;
        }
    }
    o->contrib_start[MAX_KEYS] = (uint32_t)c;
    o->contrib_wheel = cw;
    o->contrib_bus = cb;
    o->contrib_level = cl;

    memcpy(o->attack_env, t->attackEnv, sizeof(o->attack_env));
    memcpy(o->release_env, t->releaseEnv, sizeof(o->release_env));
    memcpy(o->key_comp_table, t->keyCompTable, sizeof(o->key_comp_table));
    memcpy(o->draw_bar_level, t->drawBarLevel, sizeof(o->draw_bar_level));

    o->perc_env_gain_reset_norm = t->percEnvGainResetNorm;
    o->perc_env_gain_reset_soft = t->percEnvGainResetSoft;
    o->perc_env_scaling = t->percEnvScaling;
    o->perc_decay_fast_norm = t->percEnvGainDecayFastNorm;
    o->perc_decay_fast_soft = t->percEnvGainDecayFastSoft;
    o->perc_decay_slow_norm = t->percEnvGainDecaySlowNorm;
    o->perc_decay_slow_soft = t->percEnvGainDecaySlowSoft;
    o->perc_drawbar_normal_gain = t->percDrawbarNormalGain;
    o->perc_drawbar_soft_gain = t->percDrawbarSoftGain;
    o->perc_send_bus_a = (int32_t)t->percSendBusA;
    o->perc_send_bus_b = (int32_t)t->percSendBusB;
    o->perc_trigger_bus = t->percTriggerBus;
    o->output_level_trim = t->outputLevelTrim;

    const struct b_vibrato *v = &t->inst_vibrato;
    memcpy(o->vib_offset_table[0], v->offset1Table, sizeof(v->offset1Table));
    memcpy(o->vib_offset_table[1], v->offset2Table, sizeof(v->offset2Table));
    memcpy(o->vib_offset_table[2], v->offset3Table, sizeof(v->offset3Table));
    o->vib_stator_increment = v->statorIncrement;

    memcpy(o->init_draw_bar_gain, t->drawBarGain, sizeof(o->init_draw_bar_gain));
    o->init_swell_pedal_gain = t->swellPedalGain;
    o->init_perc_env_gain_reset = t->percEnvGainReset;
    o->init_perc_env_gain_decay = t->percEnvGainDecay;
    o->init_perc_drawbar_gain = t->percDrawbarGain;
    o->init_perc_send_bus = (int32_t)t->percSendBus;
    o->init_perc_trig_restore = t->percTrigRestore;
    o->init_perc_is_soft = t->percIsSoft;
    o->init_perc_is_fast = t->percIsFast;
    o->init_perc_enabled = t->percEnabled;
    o->init_routing = t->newRouting;
    o->init_vib_table = (v->offsetTable == v->offset1Table) ? 0
                        : (v->offsetTable == v->offset2Table) ? 1 : 2;
    o->init_vib_mixed = v->mixedBuffers;
    o->init_draw_bar_change = t->drawBarChange;
    return 0;
}

void ref_free_tonegen_tables(tbf_tonegen_tables *o)
{
    free((void *)o->wave_data);
    free((void *)o->contrib_wheel);
    free((void *)o->contrib_bus);
    free((void *)o->contrib_level);
    memset(o, 0, sizeof(*o));
}

/* pointers alias the live struct b_whirl: valid until ref_destroy() */
int ref_dump_whirl_tables(void *h, tbf_whirl_tables *o)
{
    ref_inst *r = (ref_inst *)h;
    struct b_whirl *w = r->inst.whirl;
    memset(o, 0, sizeof(*o));
    o->sample_rate = w->SampleRateD;
    o->hn_fwd_displ = w->hnFwdDispl;
    o->hn_bwd_displ = w->hnBwdDispl;
    o->dr_fwd_displ = w->drFwdDispl;
    o->dr_bwd_displ = w->drBwdDispl;
    o->bfw = &w->bfw[0].b[0];
    o->bbw = &w->bbw[0].b[0];
    for (int i = 0; i < 6; ++i) {
        o->horn_phase[i] = w->hornPhase[i];
        o->drum_phase[i] = w->drumPhase[i];
        o->horn_spacing[i] = w->hornSpacing[i];
        o->drum_spacing[i] = w->drumSpacing[i];
    }
    const iir_t *src[4] = {w->hafw, w->hbfw, w->drfL, w->drfR};
    float *dst[4] = {o->hafw, o->hbfw, o->drfL, o->drfR};
    for (int k = 0; k < 4; ++k) {
        dst[k][0] = src[k][a1];
        dst[k][1] = src[k][a2];
        dst[k][2] = src[k][b0];
        dst[k][3] = src[k][b1];
        dst[k][4] = src[k][b2];
    }
    const float hap[4] = {w->haT, w->haF, w->haQ, w->haG};
    const float hbp[4] = {w->hbT, w->hbF, w->hbQ, w->hbG};
    memcpy(o->ha_param, hap, sizeof hap);
    memcpy(o->hb_param, hbp, sizeof hbp);
    o->mic_angle = w->micAngle;
    o->horn_acc = w->hornAcc;
    o->horn_dec = w->hornDec;
    o->drum_acc = w->drumAcc;
    o->drum_dec = w->drumDec;
    for (int i = 0; i < 9; ++i) {
        o->rev_horn_target[i] = w->revoptions[i].hornTarget;
        o->rev_drum_target[i] = w->revoptions[i].drumTarget;
    }
    for (int i = 0; i < 3; ++i)
        o->revselects[i] = w->revselects[i];
    o->hn_brake_pos = w->hnBrakePos;
    o->dr_brake_pos = w->drBrakePos;
    o->horn_level = w->hornLevel;
    o->leakage = w->leakage;
    o->mic_hll = w->hornMic_hll;
    o->mic_hlr = w->hornMic_hlr;
    o->mic_hrl = w->hornMic_hrl;
    o->mic_hrr = w->hornMic_hrr;
    o->mic_dll = w->drumMic_dll;
    o->mic_dlr = w->drumMic_dlr;
    o->mic_drl = w->drumMic_drl;
    o->mic_drr = w->drumMic_drr;
    o->bypass = w->bypass;
    o->init_horn_incr = w->hornIncr;
    o->init_drum_incr = w->drumIncr;
    o->init_horn_target = w->hornTarget;
    o->init_drum_target = w->drumTarget;
    o->init_horn_angle = w->hornAngleGRD;
    o->init_drum_angle = w->drumAngleGRD;
    o->init_horn_acdc = w->hornAcDc;
    o->init_drum_acdc = w->drumAcDc;
    o->init_rev_select = w->revSelect;
    return 0;
}

int ref_export_state(void *h, tbf_state_export *o)
{
    ref_inst *r = (ref_inst *)h;
    struct b_tonegen *t = r->inst.synth;
    struct b_whirl *w = r->inst.whirl;
    memset(o, 0, sizeof(*o));
    o->active_osc_l_end = t->activeOscLEnd;
    for (int i = 0; i < t->activeOscLEnd; ++i)
        o->active_osc_list[i] = t->activeOscList[i];
    for (int i = 0; i <= NOF_WHEELS; ++i) {
        o->osc_pos[i] = (uint32_t)t->oscillators[i].pos;
        o->osc_acl_pos[i] = t->oscillators[i].aclPos;
        o->ref_count[i] = t->aot[i].refCount;
    }
    o->old_routing = t->oldRouting;
    o->new_routing = t->newRouting;
    o->upper_key_count = t->upperKeyCount;
    o->key_down_count = t->keyDownCount;
    o->perc_env_gain = t->percEnvGain;
    o->key_comp_level = t->keyCompLevel;
    o->pz = t->pz;
    o->vib_stator = t->inst_vibrato.stator;
    o->vib_out_pos = t->inst_vibrato.outPos;
    memcpy(o->vib_buffer, t->inst_vibrato.vibBuffer, sizeof(o->vib_buffer));
    o->whirl_outpos = w->outpos;
    o->horn_angle = w->hornAngleGRD;
    o->drum_angle = w->drumAngleGRD;
    o->horn_incr = w->hornIncr;
    o->drum_incr = w->drumIncr;
    o->horn_acdc = w->hornAcDc;
    o->drum_acdc = w->drumAcDc;
    memcpy(o->whirl_buf[0], w->HLbuf, sizeof(w->HLbuf));
    memcpy(o->whirl_buf[1], w->HRbuf, sizeof(w->HRbuf));
    memcpy(o->whirl_buf[2], w->DLbuf, sizeof(w->DLbuf));
    memcpy(o->whirl_buf[3], w->DRbuf, sizeof(w->DRbuf));
    const iir_t *f[4] = {w->hafw, w->hbfw, w->drfL, w->drfR};
    for (int k = 0; k < 4; ++k) {
        o->iir_z[k][0] = f[k][z0];
        o->iir_z[k][1] = f[k][z1];
    }
    for (int k = 0; k < 2; ++k) {
        o->horn_iir[k][0] = f[k][a1];
        o->horn_iir[k][1] = f[k][a2];
        o->horn_iir[k][2] = f[k][b0];
        o->horn_iir[k][3] = f[k][b1];
        o->horn_iir[k][4] = f[k][b2];
    }
    o->brake_pos[0] = w->hnBrakePos;
    o->brake_pos[1] = w->drBrakePos;
    {
        const float tau[4] = {w->hornAcc, w->hornDec, w->drumAcc, w->drumDec};
        for (int k = 0; k < 4; ++k) /* as src/whirl.cpp:1255-1257 computes it per call */
            o->ramp_one_minus_l[k] = 1 - exp(-1.0 / (w->SampleRateD / (double)(size_t)128 * tau[k]));
    }
    memcpy(o->filt_z, w->z, sizeof(o->filt_z));
    memcpy(o->adx[0], w->adx0, sizeof(w->adx0));
    memcpy(o->adx[1], w->adx1, sizeof(w->adx1));
    memcpy(o->adx[2], w->adx2, sizeof(w->adx2));
    o->adi[0] = w->adi0;
    o->adi[1] = w->adi1;
    o->adi[2] = w->adi2;
    o->stat_fragments = r->stat_fragments;
    o->stat_active_wheels = r->stat_active_wheels;
    return 0;
}

unsigned ref_sizeof_tonegen_tables(void) { return (unsigned)sizeof(tbf_tonegen_tables); }
unsigned ref_sizeof_whirl_tables(void) { return (unsigned)sizeof(tbf_whirl_tables); }
unsigned ref_sizeof_state_export(void) { return (unsigned)sizeof(tbf_state_export); }

} /* extern "C" */
