/* Raw MIDI -> tbf_event compiler (include/tbf_midi.h).  Host code, no CUDA.
 *
 * Behaviour follows the reference's MIDI front end, cited per function:
 *   parse_raw_midi_data   src/midi.cpp:1261-1302
 *   process_midi_event    src/midi.cpp:1095-1255
 *   key tables            src/midi.cpp:676-726
 *   controller tables     src/midi.cpp:376-446, :591-665, :802-900, :906-1034
 *   controller functions  src/tonegen.cpp:2752-2897, src/vibrato.cpp:148-218,
 *                         src/whirl.cpp:237-261
 * The reference keeps function pointers per (channel, controller); here a function is a
 * descriptor that says which tbf_event(s) a controller value turns into.
 */
#include "tbf_midi.h"
#include "tbf_tables.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>

namespace {

enum Kind {
    K_DRAWBAR,      /* arg = bus                 setMIDIDrawBar, tonegen.cpp:2752        */
    K_PERC_ENABLE,  /*                           setPercEnableFromMIDI, :2850            */
    K_PERC_DECAY,   /*                           setPercDecayFromMIDI, :2860             */
    K_PERC_HARMONIC,/*                           setPercHarmonicFromMIDI, :2870          */
    K_PERC_VOLUME,  /*                           setPercVolumeFromMIDI, :2876            */
    K_VIB_KNOB,     /*                           setVibratoFromMIDI, vibrato.cpp:148     */
    K_VIB_ROUTING,  /*                           setVibratoRoutingFromMIDI, :177         */
    K_VIB_UPPER,    /*                           setVibratoUpperFromMIDI, :204           */
    K_VIB_LOWER,    /*                           setVibratoLowerFromMIDI, :212           */
    K_SWELL,        /*                           setSwellPedal{1,2}FromMIDI, tonegen.cpp:2885 */
    K_ROT_PRESET,   /*                           revControl, whirl.cpp:244               */
    K_ROT_TOGGLE,   /*                           setWhirlSustainPedal, whirl.cpp:252     */
    K_ROT_SELECT,   /*                           revControlAll, whirl.cpp:237            */
    K_HORN_FILTER,  /* arg = 4 * filter + {type, hz, q, gain}  setHornFilterAType ..., whirl.cpp:703-840 */
    K_WHIRL_BRAKE,  /* arg = 0 horn, 1 drum      set{Horn,Drum}BrakePosition, whirl.cpp:883 */
    K_WHIRL_RAMP,   /* arg = TBF_EV_WHIRL_RAMP's a  set{Horn,Drum}{Ac,De}celeration, :895   */
    K_UNSUPPORTED,  /* named in ccFuncNames, never registered                             */
    K_OFF_PATH      /* registered, acts on preamp / reverb only                          */
};

struct Fn {
    char name[32];
    int kind;
    int arg;
    int reg; /* position in the engine's registration sequence, 0 = never registered */
};

Fn g_fn[TBF_MIDI_MAX_FN];
int g_nfn = 0;

void add_fn(const char *name, int kind, int arg, int reg)
{
    Fn &f = g_fn[g_nfn++];
    snprintf(f.name, sizeof(f.name), "%s", name);
    f.kind = kind;
    f.arg = arg;
    f.reg = reg;
}

/* The table in ccFuncNames[] order (ids are positions; the prefix match of
 * tbf_midi_function_id depends on it) with each function's registration rank. */
void build_fn_table()
{
    if (g_nfn)
        return;
    static const char *manual[3] = {"upper", "lower", "pedal"};
    static const char *foot[9] = {"16", "513", "8", "4", "223", "2", "135", "113", "1"};
    int reg_drawbar = 3; /* after swellpedal1, swellpedal2 */
    for (int m = 0; m < 3; ++m)
        for (int d = 0; d < 9; ++d) {
            char n[32];
            snprintf(n, sizeof(n), "%s.drawbar%s", manual[m], foot[d]);
            add_fn(n, K_DRAWBAR, 9 * m + d, reg_drawbar++);
        }
    int r = reg_drawbar; /* 30 */
    add_fn("percussion.enable", K_PERC_ENABLE, 0, r + 0);
    add_fn("percussion.decay", K_PERC_DECAY, 0, r + 1);
    add_fn("percussion.harmonic", K_PERC_HARMONIC, 0, r + 2);
    add_fn("percussion.volume", K_PERC_VOLUME, 0, r + 3);
    add_fn("vibrato.knob", K_VIB_KNOB, 0, r + 4);
    add_fn("vibrato.routing", K_VIB_ROUTING, 0, r + 5);
    add_fn("vibrato.upper", K_VIB_UPPER, 0, r + 6);
    add_fn("vibrato.lower", K_VIB_LOWER, 0, r + 7);
    add_fn("swellpedal1", K_SWELL, 0, 1);
    add_fn("swellpedal2", K_SWELL, 0, 2);
    r += 8; /* preamp: 9 functions, reverb: 1, then the whirl */
    const int r_whirl = r + 10;
    add_fn("rotary.speed-preset", K_ROT_PRESET, 0, r_whirl + 1);
    add_fn("rotary.speed-toggle", K_ROT_TOGGLE, 0, r_whirl + 0);
    add_fn("rotary.speed-select", K_ROT_SELECT, 0, r_whirl + 2);
    static const char *hf[8] = {"a.type", "a.hz", "a.q", "a.gain", "b.type", "b.hz", "b.q", "b.gain"};
    for (int i = 0; i < 8; ++i) {
        char n[32];
        snprintf(n, sizeof(n), "whirl.horn.filter.%s", hf[i]);
        add_fn(n, K_HORN_FILTER, i, r_whirl + 3 + i);
    }
    static const char *df[4] = {"type", "hz", "q", "gain"};
    for (int i = 0; i < 4; ++i) {
        char n[32];
        snprintf(n, sizeof(n), "whirl.drum.filter.%s", df[i]);
        add_fn(n, K_UNSUPPORTED, 0, 0); /* named, but no module registers them */
    }
    add_fn("whirl.horn.brakepos", K_WHIRL_BRAKE, 0, r_whirl + 11);
    add_fn("whirl.drum.brakepos", K_WHIRL_BRAKE, 1, r_whirl + 12);
    add_fn("whirl.horn.acceleration", K_WHIRL_RAMP, 0, r_whirl + 13);
    add_fn("whirl.horn.deceleration", K_WHIRL_RAMP, 1, r_whirl + 14);
    add_fn("whirl.drum.acceleration", K_WHIRL_RAMP, 2, r_whirl + 15);
    add_fn("whirl.drum.deceleration", K_WHIRL_RAMP, 3, r_whirl + 16);
    /* preamp registers xov.ctl_biased, _fb, _fb2, _gfb, sagtobias, character, enable,
     * inputgain, outputgain in that order (src/overdrive.cpp:588-598) */
    add_fn("overdrive.enable", K_OFF_PATH, 0, r + 6);
    add_fn("overdrive.character", K_OFF_PATH, 0, r + 5);
    add_fn("overdrive.inputgain", K_OFF_PATH, 0, r + 7);
    add_fn("overdrive.outputgain", K_OFF_PATH, 0, r + 8);
    add_fn("xov.ctl_biased_fb2", K_OFF_PATH, 0, r + 2);
    add_fn("xov.ctl_biased_fb", K_OFF_PATH, 0, r + 1);
    add_fn("xov.ctl_biased_gfb", K_OFF_PATH, 0, r + 3);
    add_fn("xov.ctl_biased", K_OFF_PATH, 0, r + 0);
    add_fn("xov.ctl_sagtobias", K_OFF_PATH, 0, r + 4);
    add_fn("reverb.mix", K_OFF_PATH, 0, r + 9);
    add_fn("convolution.mix", K_OFF_PATH, 0, 0); /* only with the convolver compiled in */
}

int fn_id(const char *name)
{
    build_fn_table();
    /* first table entry that is a prefix of `name` (getCCFunctionId, midi.cpp:333-346) */
    for (int i = 0; i < g_nfn; ++i)
        if (strncmp(name, g_fn[i].name, strlen(g_fn[i].name)) == 0)
            return i;
    return -1;
}

/* loadCCMap (midi.cpp:802-826): one controller per function and channel set, last call wins */
void load_cc(tbf_midi_map *m, const char *name, int cc, bool a, bool b, bool c)
{
    const int x = fn_id(name);
    if (x < 0)
        return;
    if (a)
        m->cc_use[0][x] = (uint8_t)cc;
    if (b)
        m->cc_use[1][x] = (uint8_t)cc;
    if (c)
        m->cc_use[2][x] = (uint8_t)cc;
}

/* remove_CC_map (midi.cpp:625-659).  It clears ctrlflg[channel][FUNCTION index], not
 * [controller]: kept as the reference has it */
void remove_cc(tbf_midi_map *m, int set, int cc)
{
    const int ch = m->rcv_ch[set];
    for (int i = 0; i < 127 && i < TBF_MIDI_MAX_FN; ++i)
        if (m->cc_use[set][i] == cc) {
            m->cc_use[set][i] = TBF_MIDI_UNMAPPED;
            m->cc_flags[ch][i] = 0;
            return;
        }
}

bool parse_int(const char *s, int lo, int hi, int *out)
{
    if (!s)
        return false;
    char *end = NULL;
    const long v = strtol(s, &end, 10);
    if (end == s)
        return false;
    while (*end == ' ' || *end == '\t' || *end == '\r' || *end == '\n')
        ++end;
    if (*end || v < lo || v > hi)
        return false;
    *out = (int)v;
    return true;
}

inline tbf_event make_event(uint32_t instance, uint32_t fragment, int type, int a, int b)
{
    tbf_event e;
    e.instance = instance;
    e.fragment = fragment;
    e.type = type;
    e.a = a;
    e.b = b;
    return e;
}

/* loadKeyTableA / B / C (midi.cpp:676-726) for one note of channel set `set` */
struct KeyInts {
    int transpose, nsh_a, nsh_a_u, nsh_a_pl, nsh_a_ul, nsh_b, nsh_c, split_a_pl, split_a_ul;
};
inline int key_of(const KeyInts &k, int set, int i)
{
    if (set == 1)
        return 128 + i + k.transpose + k.nsh_b;
    if (set == 2)
        return 2 * 128 + i + k.transpose + k.nsh_c;
    int off;
    if (i < k.split_a_pl)
        off = k.transpose + k.nsh_a_pl + 2 * 128;
    else if (i < k.split_a_ul)
        off = k.transpose + k.nsh_a_ul + 128;
    else if (k.split_a_ul)
        off = k.transpose + k.nsh_a_u;
    else
        off = k.transpose + k.nsh_a;
    return i + off;
}

/* revSelect after useRevOption(w, n, 2) (whirl.cpp:205-218) */
inline int rev_select_of_option(int n)
{
    const int hr = (n / 3) % 3;
    return hr == 2 ? 2 /* WHIRL_FAST */ : (hr == 1 ? 0 /* WHIRL_SLOW */ : 1 /* WHIRL_STOP */);
}

/* installProgram (src/program.cpp:735-921): the engine calls a programme change stands for,
 * in the reference's order.  The setters it reaches through callMIDIControlFunction are the
 * registered MIDI functions, so their value quantisation applies (knob * 23 -> knob, ...). */
int install_program(const tbf_midi_map *m, tbf_midi_state *s, uint32_t instance, uint32_t fragment,
                    int number, tbf_event *out, size_t cap)
{
    const tbf_programme_table *t = m->programmes;
    if (!t) {
        s->n_unsupported++;
        return 0;
    }
    const int p = number + t->controller_offset;
    if (!(0 < p && p < TBF_MAXPROGS)) {
        s->n_ignored++;
        return 0;
    }
    const tbf_programme &g = t->prog[p];
    const uint32_t f = g.flags;
    if (!(f & TBF_PGM_INUSE)) {
        s->n_ignored++;
        return 0;
    }
    size_t n = 0;
#define PUT(type, a, b)                                                                               \
    do {                                                                                              \
        if (n >= cap)                                                                                 \
            return TBF_ERR_ARG;                                                                       \
        out[n++] = make_event(instance, fragment, (type), (a), (b));                                  \
        if ((type) == TBF_EV_VIB_UPPER)                                                               \
            s->vib_routing = (s->vib_routing & ~2) | ((a) ? 2 : 0);                                   \
        if ((type) == TBF_EV_VIB_LOWER)                                                               \
            s->vib_routing = (s->vib_routing & ~1) | ((a) ? 1 : 0);                                   \
    } while (0)

    if (f & TBF_PGM_DRWRND) /* randomizeDrawbars() draws from the process-global rand() */
        s->n_unsupported++;
    static const uint32_t manual_flag[3] = {TBF_PGM_DRAWBR, TBF_PGM_LOWDRW, TBF_PGM_PDLDRW};
    for (int man = 0; man < 3; ++man) /* setDrawBars, tonegen.cpp:2769-2796 */
        if ((f & manual_flag[man]) && !(f & TBF_PGM_DRWRND))
            for (int i = 0; i < 9; ++i)
                PUT(TBF_EV_DRAWBAR, 9 * man + i, g.drawbars[man][i]);
    if (f & TBF_PGM_SCANNR) {
        const int knob = ((g.scanner & 0xf) << 1) - ((g.scanner & 0x80) ? 1 : 2);
        if (knob >= 0 && knob <= 5)
            PUT(TBF_EV_VIB_KNOB, knob, 0);
    }
    if (f & TBF_PGM_VCRUPR) { /* vibrato.routing(rt << 5): both manuals are written */
        const int rt = (s->vib_routing & ~2) | ((g.scanner & 0x200) ? 2 : 0);
        PUT(TBF_EV_VIB_UPPER, rt >> 1, 0);
        PUT(TBF_EV_VIB_LOWER, rt & 1, 0);
    }
    if (f & TBF_PGM_VCRLWR) {
        const int rt = (s->vib_routing & ~1) | ((g.scanner & 0x100) ? 1 : 0);
        PUT(TBF_EV_VIB_UPPER, rt >> 1, 0);
        PUT(TBF_EV_VIB_LOWER, rt & 1, 0);
    }
    if (f & TBF_PGM_PRCENA) { /* the setter, then the registered function: two calls */
        PUT(TBF_EV_PERC_ENABLE, g.perc_enabled ? 1 : 0, 0);
        PUT(TBF_EV_PERC_ENABLE, g.perc_enabled ? 1 : 0, 0);
    }
    if (f & TBF_PGM_PRCVOL)
        PUT(TBF_EV_PERC_SOFT, g.perc_soft ? 1 : 0, 0);
    if (f & TBF_PGM_PRCSPD)
        PUT(TBF_EV_PERC_FAST, g.perc_fast ? 1 : 0, 0);
    if (f & TBF_PGM_PRCHRM)
        PUT(TBF_EV_PERC_FIRST, g.perc_second ? 1 : 0, 0);
    if (f & TBF_PGM_ROTSPS) { /* rotary.speed-preset(ceilf(sel * 63.5f)) -> setRevSelect(u / 43) */
        const int u = (int)ceilf(g.rotary_speed * 63.5f);
        PUT(TBF_EV_REV_SELECT, u / 43, 0);
        s->rev_select = (u / 43) % 3;
    }
    /* keyboard splits and transposes change the instance's note -> key mapping */
    if (f & TBF_PGM_KSPLTP)
        s->split_a_pl = g.split_pedals;
    if (f & TBF_PGM_KSPLTL)
        s->split_a_ul = g.split_lower;
    if (f & TBF_PGM_TRA_PD)
        s->nsh_a_pl = g.transpose[5];
    if (f & TBF_PGM_TRA_LM)
        s->nsh_a_ul = g.transpose[6];
    if (f & TBF_PGM_TRA_UM)
        s->nsh_a_u = g.transpose[4];
    if (f & TBF_PGM_TRANSP)
        s->transpose = g.transpose[0];
    if (f & TBF_PGM_TRCH_A)
        s->nsh_a = g.transpose[1];
    if (f & TBF_PGM_TRCH_B)
        s->nsh_b = g.transpose[2];
    if (f & TBF_PGM_TRCH_C)
        s->nsh_c = g.transpose[3];
#undef PUT
    return (int)n;
}

bool affirmative(const char *v) /* isAffirmative, program.cpp:242-259 */
{
    int n;
    if (!strcasecmp(v, "on") || !strcasecmp(v, "yes") || !strcasecmp(v, "true") ||
        !strcasecmp(v, "enabled"))
        return true;
    return sscanf(v, "%d", &n) == 1 && n != 0;
}

bool negatory(const char *v) /* isNegatory, program.cpp:264-283 */
{
    int n;
    if (!strcasecmp(v, "off") || !strcasecmp(v, "no") || !strcasecmp(v, "none") ||
        !strcasecmp(v, "false") || !strcasecmp(v, "disabled"))
        return true;
    return sscanf(v, "%d", &n) == 1 && n == 0;
}

bool any_of(const char *v, const char *a, const char *b = NULL, const char *c = NULL,
            const char *d = NULL)
{
    const char *opts[4] = {a, b, c, d};
    for (int i = 0; i < 4; ++i)
        if (opts[i] && !strcasecmp(v, opts[i]))
            return true;
    return false;
}

/* parseDrawbarRegistration, program.cpp:203-237: nine digits 0..8, blanks, '-' and '_' skipped */
bool parse_registration(const char *v, uint8_t bar[9])
{
    int bus = 0;
    for (const char *t = v; bus < 9; ++t) {
        if (*t == '\0')
            return false;
        if (*t == ' ' || *t == '\t' || *t == '\n' || *t == '\r' || *t == '\f' || *t == '\v' ||
            *t == '-' || *t == '_')
            continue;
        if (*t < '0' || *t > '8')
            return false;
        bar[bus++] = (uint8_t)(*t - '0');
    }
    return true;
}

} // namespace

extern "C" {

int tbf_midi_function_count(void)
{
    build_fn_table();
    return g_nfn;
}

const char *tbf_midi_function_name(int id)
{
    build_fn_table();
    return (id >= 0 && id < g_nfn) ? g_fn[id].name : NULL;
}

int tbf_midi_function_id(const char *name) { return name ? fn_id(name) : -1; }

void tbf_midi_map_default(tbf_midi_map *m)
{
    if (!m)
        return;
    build_fn_table();
    memset(m, 0, sizeof(*m));
    m->rcv_ch[0] = 0;
    m->rcv_ch[1] = 1;
    m->rcv_ch[2] = 2;
    memset(m->cc_use, TBF_MIDI_UNMAPPED, sizeof(m->cc_use));
    m->revselects[0] = 4; /* computeRotationSpeeds, whirl.cpp:289-291 */
    m->revselects[1] = 0;
    m->revselects[2] = 8;
    /* midiPrimeControllerMapping, midi.cpp:839-900 */
    load_cc(m, "swellpedal1", 7, true, true, true);
    load_cc(m, "swellpedal2", 11, true, true, true);
    static const char *manual[3] = {"upper", "lower", "pedal"};
    static const char *foot[9] = {"16", "513", "8", "4", "223", "2", "135", "113", "1"};
    for (int s = 0; s < 3; ++s)
        for (int d = 0; d < 9; ++d) {
            char n[32];
            snprintf(n, sizeof(n), "%s.drawbar%s", manual[s], foot[d]);
            load_cc(m, n, 70 + d, s == 0, s == 1, s == 2);
        }
    load_cc(m, "rotary.speed-preset", 1, true, true, true);
    load_cc(m, "rotary.speed-toggle", 64, true, true, true);
    load_cc(m, "reverb.mix", 91, true, false, false);
    load_cc(m, "percussion.enable", 66, true, true, true);
    load_cc(m, "percussion.enable", 80, true, false, false);
    load_cc(m, "percussion.volume", 81, true, false, false);
    load_cc(m, "percussion.decay", 82, true, false, false);
    load_cc(m, "percussion.harmonic", 83, true, false, false);
    load_cc(m, "vibrato.routing", 95, true, false, false);
    load_cc(m, "vibrato.knob", 92, true, false, false);
    load_cc(m, "vibrato.lower", 30, true, false, false);
    load_cc(m, "vibrato.upper", 31, true, false, false);
    load_cc(m, "overdrive.enable", 65, true, true, true);
    load_cc(m, "overdrive.character", 93, true, false, false);
    load_cc(m, "convolution.mix", 94, true, false, false);
    tbf_midi_map_finish(m, NULL);
}

int tbf_midi_map_apply_cfg(tbf_midi_map *m, const char *key, const char *value)
{
    if (!m || !key)
        return TBF_ERR_ARG;
    if (strncasecmp(key, "midi.", 5) != 0)
        return 0;
    struct {
        const char *name;
        int lo, hi;
    } ints[] = {{"midi.upper.channel", 1, 16},
                {"midi.lower.channel", 1, 16},
                {"midi.pedals.channel", 1, 16},
                {"midi.transpose", -127, 127},
                {"midi.upper.transpose", -127, 127},
                {"midi.lower.transpose", -127, 127},
                {"midi.pedals.transpose", -127, 127},
                {"midi.pedals.transpose.split", -127, 127},
                {"midi.lower.transpose.split", -127, 127},
                {"midi.upper.transpose.split", -127, 127}};
    for (unsigned i = 0; i < sizeof(ints) / sizeof(ints[0]); ++i) {
        if (strcasecmp(key, ints[i].name) != 0)
            continue;
        int v;
        if (!parse_int(value, ints[i].lo, ints[i].hi, &v))
            return TBF_ERR_ARG;
        switch (i) {
        case 0: m->rcv_ch[0] = (uint8_t)(v - 1); break;
        case 1: m->rcv_ch[1] = (uint8_t)(v - 1); break;
        case 2: m->rcv_ch[2] = (uint8_t)(v - 1); break;
        case 3: m->transpose = v; break;
        case 4: m->nsh_a = v; break;
        case 5: m->nsh_b = v; break;
        case 6: m->nsh_c = v; break;
        case 7: m->nsh_a_pl = v; break;
        case 8: m->nsh_a_ul = v; break;
        default: m->nsh_a_u = v; break;
        }
        return 1;
    }
    if (strncasecmp(key, "midi.controller.reset", 21) == 0) {
        if (!value || !*value)
            return 0;
        if (atoi(value)) { /* clearControllerMapping, midi.cpp:591-623 */
            memset(m->cc_use, TBF_MIDI_UNMAPPED, sizeof(m->cc_use));
            memset(m->cc_flags, 0, sizeof(m->cc_flags));
        }
        return 1;
    }
    if (strncasecmp(key, "midi.controller.", 16) == 0) {
        int set, idx;
        if (strncasecmp(key + 16, "upper", 5) == 0) {
            set = 0;
            idx = 22;
        } else if (strncasecmp(key + 16, "lower", 5) == 0) {
            set = 1;
            idx = 22;
        } else if (strncasecmp(key + 16, "pedals", 6) == 0) {
            set = 2;
            idx = 23;
        } else {
            return TBF_ERR_ARG;
        }
        int cc;
        if (strlen(key) <= (size_t)idx || sscanf(key + idx, "%d", &cc) != 1)
            return 0;
        if (cc < 0 || cc >= 128 || !value)
            return TBF_ERR_ARG;
        const int ch = m->rcv_ch[set];
        if (strcmp(value, "unmap") == 0) {
            remove_cc(m, set, cc);
            return 0; /* the reference does not count an unmap as consumed */
        }
        const int x = fn_id(value);
        if (x < 0)
            return TBF_ERR_ARG;
        remove_cc(m, set, cc);
        m->cc_use[set][x] = (uint8_t)cc;
        const size_t l = strlen(value);
        if (l && value[l - 1] == '-') /* parseCCFlags, midi.cpp:348-353 */
            m->cc_flags[ch][cc] |= TBF_MIDI_FLAG_INV;
        return 1;
    }
    return 0;
}

void tbf_midi_map_split(tbf_midi_map *m, int flags, int split_a_pl, int split_a_ul, int nsh_a_pl,
                        int nsh_a_ul, int nsh_a_u)
{
    if (!m)
        return;
    if (flags & 1)
        m->split_a_pl = split_a_pl;
    if (flags & 2)
        m->split_a_ul = split_a_ul;
    if (flags & 4)
        m->nsh_a_pl = nsh_a_pl;
    if (flags & 8)
        m->nsh_a_ul = nsh_a_ul;
    if (flags & 16)
        m->nsh_a_u = nsh_a_u;
}

int tbf_midi_map_finish(tbf_midi_map *m, const int32_t revselects[3])
{
    if (!m)
        return TBF_ERR_ARG;
    build_fn_table();
    if (revselects)
        memcpy(m->revselects, revselects, sizeof(m->revselects));
    /* key tables A, B, C (midi.cpp:676-726) */
    const KeyInts ki = {m->transpose, m->nsh_a, m->nsh_a_u, m->nsh_a_pl, m->nsh_a_ul,
                        m->nsh_b,     m->nsh_c, m->split_a_pl, m->split_a_ul};
    int16_t kt[3][128];
    for (int set = 0; set < 3; ++set)
        for (int i = 0; i < 128; ++i)
            kt[set][i] = (int16_t)key_of(ki, set, i);
    /* controller vectors A, B, C: the engine modules register their functions one by one,
     * each taking the controller its id is mapped to (midi.cpp:416-446) */
    int16_t vec[3][128];
    memset(vec, 0xFF, sizeof(vec));
    int order[TBF_MIDI_MAX_FN], n = 0;
    for (int rank = 1; rank <= TBF_MIDI_MAX_FN; ++rank)
        for (int i = 0; i < g_nfn; ++i)
            if (g_fn[i].reg == rank)
                order[n++] = i;
    for (int k = 0; k < n; ++k)
        for (int s = 0; s < 3; ++s) {
            const int cc = m->cc_use[s][order[k]];
            if (cc < 128)
                vec[s][cc] = (int16_t)order[k];
        }
    /* status table (midi.cpp:782-800): a channel shared by two manuals goes to the later */
    memset(m->key_table, 0xFF, sizeof(m->key_table));
    memset(m->cc_fn, 0xFF, sizeof(m->cc_fn));
    memset(m->ch_set, 0xFF, sizeof(m->ch_set));
    for (int s = 0; s < 3; ++s) {
        const int ch = m->rcv_ch[s] & 15;
        memcpy(m->key_table[ch], kt[s], sizeof(kt[s]));
        memcpy(m->cc_fn[ch], vec[s], sizeof(vec[s]));
        m->ch_set[ch] = (int8_t)s;
    }
    return TBF_OK;
}

void tbf_midi_state_init(tbf_midi_state *s, const tbf_midi_map *m, const tbf_whirl_tables *w)
{
    if (!s)
        return;
    memset(s, 0, sizeof(*s));
    if (w) {
        s->rev_select = w->init_rev_select % 3;
        memcpy(s->horn_filter[0], w->ha_param, sizeof(w->ha_param));
        memcpy(s->horn_filter[1], w->hb_param, sizeof(w->hb_param));
        s->sample_rate = w->sample_rate;
    } else {
        tbf_whirl_spec d;
        tbf_whirl_spec_default(&d, 48000.0);
        const float a[4] = {(float)d.ha_type, d.ha_f, d.ha_q, d.ha_g};
        const float b[4] = {(float)d.hb_type, d.hb_f, d.hb_q, d.hb_g};
        memcpy(s->horn_filter[0], a, sizeof a);
        memcpy(s->horn_filter[1], b, sizeof b);
        s->sample_rate = 48000.0;
    }
    if (m) {
        s->transpose = m->transpose;
        s->nsh_a = m->nsh_a;
        s->nsh_a_u = m->nsh_a_u;
        s->nsh_a_pl = m->nsh_a_pl;
        s->nsh_a_ul = m->nsh_a_ul;
        s->nsh_b = m->nsh_b;
        s->nsh_c = m->nsh_c;
        s->split_a_pl = m->split_a_pl;
        s->split_a_ul = m->split_a_ul;
    }
}

int tbf_midi_key(const tbf_midi_map *m, const tbf_midi_state *s, int channel, int note)
{
    if (!m || !s || channel < 0 || channel > 15 || note < 0 || note > 127)
        return -1;
    const int set = m->ch_set[channel];
    if (set < 0)
        return -1;
    const KeyInts ki = {s->transpose, s->nsh_a, s->nsh_a_u, s->nsh_a_pl, s->nsh_a_ul,
                        s->nsh_b,     s->nsh_c, s->split_a_pl, s->split_a_ul};
    const int key = key_of(ki, set, note);
    /* the reference keeps keys in a `short` table and plays only non-negative ones */
    return key < 0 ? -1 : key;
}

int tbf_midi_compile(const tbf_midi_map *m, tbf_midi_state *s, uint32_t instance, uint32_t fragment,
                     const uint8_t *msg, size_t size, tbf_event *out, size_t cap)
{
    if (!m || !s || !msg || (!out && cap))
        return TBF_ERR_ARG;
    build_fn_table();
    /* parse_raw_midi_data: 3-byte messages, 2 bytes only for a program change */
    if (size < 2 || size > 3 || (size == 2 && (msg[0] & 0xF0) != 0xC0)) {
        s->n_ignored++;
        return 0;
    }
    const int ch = msg[0] & 0x0F;
    const int status = msg[0] & 0xF0;
    size_t n = 0;
#define EMIT(type, a, b)                                                                              \
    do {                                                                                              \
        if (n >= cap)                                                                                 \
            return TBF_ERR_ARG;                                                                       \
        out[n++] = make_event(instance, fragment, (type), (a), (b));                                  \
        if ((type) == TBF_EV_VIB_UPPER)                                                               \
            s->vib_routing = (s->vib_routing & ~2) | ((a) ? 2 : 0);                                   \
        if ((type) == TBF_EV_VIB_LOWER)                                                               \
            s->vib_routing = (s->vib_routing & ~1) | ((a) ? 1 : 0);                                   \
    } while (0)

    if (status == 0x80 || status == 0x90) {
        const int note = msg[1] & 0x7F, vel = msg[2] & 0x7F;
        const int key = tbf_midi_key(m, s, ch, note);
        if (key < 0) { /* no table on this channel, or transposed below key 0 */
            s->n_ignored++;
            return 0;
        }
        if (key >= TBF_MAX_KEYS) {
            s->n_unmapped++;
            return 0;
        }
        EMIT((status == 0x90 && vel > 0) ? TBF_EV_KEY_ON : TBF_EV_KEY_OFF, key, 0);
        return (int)n;
    }
    if (status == 0xC0) {
        const int rc = install_program(m, s, instance, fragment, msg[1] & 0x7F, out, cap);
        return rc;
    }
    if (status != 0xB0) { /* pitch bend and everything else */
        s->n_ignored++;
        return 0;
    }
    const int param = msg[1] & 0x7F;
    int val = msg[2] & 0x7F;
    if (param == 0x00 || param == 0x20 || param == 121) {
        s->n_ignored++;
        return 0;
    }
    if (param == 120 || param == 123) { /* midi_panic, midi.cpp:1085-1093 */
        for (int k = 0; k < TBF_MAX_KEYS; ++k)
            EMIT(TBF_EV_KEY_OFF, k, 0);
        return (int)n;
    }
    if (param >= 120) {
        s->n_ignored++;
        return 0;
    }
    const int f = m->cc_fn[ch][param];
    if (f < 0) {
        s->n_ignored++;
        return 0;
    }
    if (m->cc_flags[ch][param] & TBF_MIDI_FLAG_INV)
        val = 127 - val;
    const Fn &fn = g_fn[f];
    switch (fn.kind) {
    case K_DRAWBAR:
        EMIT(TBF_EV_DRAWBAR, fn.arg, (int)rint((127 - val) * 8.0 / 127.0));
        break;
    case K_PERC_ENABLE:
        EMIT(TBF_EV_PERC_ENABLE, val < 64 ? 0 : 1, 0);
        break;
    case K_PERC_DECAY:
        EMIT(TBF_EV_PERC_FAST, val < 64 ? 0 : 1, 0);
        break;
    case K_PERC_HARMONIC:
        EMIT(TBF_EV_PERC_FIRST, val < 64 ? 0 : 1, 0);
        break;
    case K_PERC_VOLUME:
        EMIT(TBF_EV_PERC_SOFT, val < 64 ? 0 : 1, 0);
        break;
    case K_VIB_KNOB:
        if (val / 23 <= 5)
            EMIT(TBF_EV_VIB_KNOB, val / 23, 0);
        break;
    case K_VIB_ROUTING:
        EMIT(TBF_EV_VIB_UPPER, (val / 32) >> 1, 0);
        EMIT(TBF_EV_VIB_LOWER, (val / 32) & 1, 0);
        break;
    case K_VIB_UPPER:
        EMIT(TBF_EV_VIB_UPPER, val < 64 ? 0 : 1, 0);
        break;
    case K_VIB_LOWER:
        EMIT(TBF_EV_VIB_LOWER, val < 64 ? 0 : 1, 0);
        break;
    case K_SWELL:
        EMIT(TBF_EV_SWELL, val, 0);
        break;
    case K_ROT_PRESET: /* setRevSelect(w, u / 43): revSelect = n % 3, option by table */
        EMIT(TBF_EV_REV_SELECT, val / 43, 0);
        s->rev_select = (val / 43) % 3;
        break;
    case K_ROT_SELECT:
        EMIT(TBF_EV_REV_OPTION, val / 15, 0);
        s->rev_select = rev_select_of_option(val / 15);
        break;
    case K_ROT_TOGGLE:
        if (val > 63) {
            const int opt = (s->rev_select == 0) ? m->revselects[2] : m->revselects[0];
            EMIT(TBF_EV_REV_OPTION, opt, 0);
            s->rev_select = rev_select_of_option(opt);
        }
        break;
    case K_HORN_FILTER: {
        /* setHornFilterAType ... setHornFilterBGain (src/whirl.cpp:703-840): the value in
         * double, the field a float, then UPDATE_x_FILTER = setIIRFilter on all four */
        float *q = s->horn_filter[fn.arg >> 2];
        const double u = (double)val;
        switch (fn.arg & 3) {
        case 0:
            q[0] = (float)(int)(val / 15);
            break;
        case 1:
            q[1] = (float)(250.0 + ((8000.0 - 250.0) * ((u * u) / 16129.0)));
            break;
        case 2:
            q[2] = (float)(0.01 + ((6.00 - 0.01) * (u / 127.0)));
            break;
        default:
            q[3] = (float)(-48.0 + ((48.0 - -48.0) * (u / 127.0)));
            break;
        }
        float c5[5];
        if (tbf_iir_coefficients(c5, (int)q[0], q[1], q[2], q[3], s->sample_rate) == 1) {
            for (int k = 0; k < 5; ++k) {
                int32_t bits;
                memcpy(&bits, &c5[k], sizeof bits);
                EMIT(TBF_EV_HORN_FILTER, 8 * (fn.arg >> 2) + k, bits);
            }
        } /* else: out of setIIRFilter's range, the old coefficients stay */
        break;
    }
    case K_WHIRL_BRAKE:
        EMIT(TBF_EV_WHIRL_BRAKE, fn.arg, val);
        break;
    case K_WHIRL_RAMP:
        EMIT(TBF_EV_WHIRL_RAMP, fn.arg, val);
        break;
    case K_UNSUPPORTED:
        s->n_unsupported++;
        break;
    default:
        s->n_ignored++;
        break;
    }
#undef EMIT
    return (int)n;
}

void tbf_programmes_init(tbf_programme_table *t)
{
    if (!t)
        return;
    memset(t, 0, sizeof(*t));
    t->controller_offset = 1;
    t->previous = -1;
}

int tbf_programme_bind(tbf_programme_table *t, int pgmnr, const char *sym, const char *val)
{
    if (!t || !sym || !val || pgmnr < 0 || pgmnr >= TBF_MAXPROGS)
        return TBF_ERR_ARG;
    tbf_programme &g = t->prog[pgmnr];
    if (pgmnr != t->previous) { /* a new programme number starts from no properties */
        g.flags = 0;
        t->previous = pgmnr;
    }
    const uint32_t U = TBF_PGM_INUSE;
    int iv;
    float fv;
    /* keywords: propertySymbols[], program.cpp:132-163 */
    if (!strcasecmp(sym, "name")) {
        snprintf(g.name, sizeof(g.name), "%s", val);
        g.flags |= U;
    } else if (!strcasecmp(sym, "drawbars") || !strcasecmp(sym, "drawbarsupper") ||
               !strcasecmp(sym, "drawbarslower") || !strcasecmp(sym, "drawbarspedals")) {
        const int man = !strcasecmp(sym, "drawbarslower") ? 1 : (!strcasecmp(sym, "drawbarspedals") ? 2 : 0);
        const uint32_t fl = man == 0 ? TBF_PGM_DRAWBR : (man == 1 ? TBF_PGM_LOWDRW : TBF_PGM_PDLDRW);
        if (!strcasecmp(val, "random"))
            g.flags |= U | fl | TBF_PGM_DRWRND;
        else if (parse_registration(val, g.drawbars[man]))
            g.flags |= U | fl;
        else
            return TBF_ERR_ARG;
    } else if (!strcasecmp(sym, "attackenv") || !strcasecmp(sym, "attacklvl") ||
               !strcasecmp(sym, "attackdur")) {
        /* accepted and unused, as in the reference */
    } else if (!strcasecmp(sym, "vibrato") || !strcasecmp(sym, "vibratoknob")) {
        static const char *names[6] = {"v1", "v2", "v3", "c1", "c2", "c3"};
        static const int codes[6] = {0x01, 0x02, 0x03, 0x81, 0x82, 0x83};
        int k = -1;
        for (int i = 0; i < 6; ++i)
            if (!strcasecmp(val, names[i]))
                k = i;
        if (k < 0)
            return TBF_ERR_ARG;
        g.scanner = (int16_t)((g.scanner & 0xFF00) | codes[k]);
        g.flags |= U | TBF_PGM_SCANNR;
    } else if (!strcasecmp(sym, "vibratoupper") || !strcasecmp(sym, "vibratolower")) {
        const int bit = !strcasecmp(sym, "vibratoupper") ? 0x200 : 0x100;
        const uint32_t fl = bit == 0x200 ? TBF_PGM_VCRUPR : TBF_PGM_VCRLWR;
        if (negatory(val))
            g.scanner = (int16_t)(g.scanner & ~bit);
        else if (affirmative(val))
            g.scanner = (int16_t)(g.scanner | bit);
        else
            return TBF_ERR_ARG;
        g.flags |= U | fl;
    } else if (!strcasecmp(sym, "perc")) {
        if (affirmative(val))
            g.perc_enabled = 1;
        else if (negatory(val))
            g.perc_enabled = 0;
        else
            return TBF_ERR_ARG;
        g.flags |= U | TBF_PGM_PRCENA;
    } else if (!strcasecmp(sym, "percvol")) {
        if (any_of(val, "normal", "high", "hi"))
            g.perc_soft = 0;
        else if (any_of(val, "soft", "low", "lo"))
            g.perc_soft = 1;
        else
            return TBF_ERR_ARG;
        g.flags |= U | TBF_PGM_PRCVOL;
    } else if (!strcasecmp(sym, "percspeed")) {
        if (any_of(val, "fast", "high", "hi"))
            g.perc_fast = 1;
        else if (any_of(val, "slow", "low", "lo"))
            g.perc_fast = 0;
        else
            return TBF_ERR_ARG;
        g.flags |= U | TBF_PGM_PRCSPD;
    } else if (!strcasecmp(sym, "percharm")) {
        if (any_of(val, "second", "2nd", "low", "lo"))
            g.perc_second = 1;
        else if (any_of(val, "third", "3rd", "high", "hi"))
            g.perc_second = 0;
        else
            return TBF_ERR_ARG;
        g.flags |= U | TBF_PGM_PRCHRM;
    } else if (!strcasecmp(sym, "overdrive")) {
        if (negatory(val))
            g.overdrive = 0;
        else if (affirmative(val))
            g.overdrive = 1;
        else
            return TBF_ERR_ARG;
        g.flags |= U | TBF_PGM_OVRSEL;
    } else if (!strcasecmp(sym, "rotary")) {
        /* pr_RotaryEnabled has no case in bindToProgram: accepted, no effect */
    } else if (!strcasecmp(sym, "rotaryspeed")) {
        if (any_of(val, "tremolo", "fast", "high", "hi"))
            g.rotary_speed = 2;
        else if (any_of(val, "chorale", "slow", "low", "lo"))
            g.rotary_speed = 0;
        else if (any_of(val, "stop", "zero", "break", "stopped"))
            g.rotary_speed = 1;
        else
            return TBF_ERR_ARG;
        g.flags |= U | TBF_PGM_ROTSPS;
    } else if (!strcasecmp(sym, "reverbmix")) {
        g.flags |= U | TBF_PGM_RVBMIX;
        if (sscanf(val, "%f", &fv) != 1 || fv < 0.0f || fv > 1.0f)
            return TBF_ERR_ARG;
        g.reverb_mix = fv;
    } else if (!strcasecmp(sym, "keysplitlower") || !strcasecmp(sym, "keysplitpedals")) {
        const bool lower = !strcasecmp(sym, "keysplitlower");
        g.flags |= U | (lower ? TBF_PGM_KSPLTL : TBF_PGM_KSPLTP);
        if (sscanf(val, "%d", &iv) != 1 || iv < 0 || iv > 127)
            return TBF_ERR_ARG;
        (lower ? g.split_lower : g.split_pedals) = (int16_t)iv;
    } else {
        static const struct {
            const char *name;
            uint32_t flag;
            int slot;
        } tr[7] = {{"transpose", TBF_PGM_TRANSP, 0},        {"transposeupper", TBF_PGM_TRCH_A, 1},
                   {"transposelower", TBF_PGM_TRCH_B, 2},   {"transposepedals", TBF_PGM_TRCH_C, 3},
                   {"trssplitupper", TBF_PGM_TRA_UM, 4},    {"trssplitpedals", TBF_PGM_TRA_PD, 5},
                   {"trssplitlower", TBF_PGM_TRA_LM, 6}};
        int k = -1;
        for (int i = 0; i < 7; ++i)
            if (!strcasecmp(sym, tr[i].name))
                k = i;
        if (k < 0)
            return TBF_ERR_ARG; /* unrecognised property */
        g.flags |= U | tr[k].flag;
        if (sscanf(val, "%d", &iv) != 1 || iv < -127 || iv > 127)
            return TBF_ERR_ARG;
        g.transpose[tr[k].slot] = (int16_t)iv;
    }
    return TBF_OK;
}

int tbf_programmes_load(tbf_programme_table *t, const char *text, int *err_line)
{
    if (!t || !text)
        return TBF_ERR_ARG;
    /* tokens (getToken, pgmParser.cpp:80-196): '{' '}' '=' ',' and strings -- quoted with
     * backslash escapes, or bare runs of [A-Za-z0-9-._+]; '#' starts a comment */
    const char *p = text;
    int line = 1;
    enum { T_EOF = 0, T_STR = 1, T_ERR = 2 };
    char tok[256];
    auto next = [&]() -> int {
        for (;;) {
            while (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n' || *p == '\f' || *p == '\v') {
                if (*p == '\n')
                    ++line;
                ++p;
            }
            if (*p != '#')
                break;
            while (*p && *p != '\n')
                ++p;
        }
        if (!*p)
            return T_EOF;
        size_t n = 0;
        tok[0] = 0;
        if (*p == '{' || *p == '}' || *p == '=' || *p == ',')
            return *p++;
        if (*p == '"') {
            ++p;
            for (;;) {
                if (!*p)
                    return T_ERR;
                if (*p == '"') {
                    ++p;
                    break;
                }
                if (*p == '\\' && p[1])
                    ++p;
                if (*p == '\n')
                    ++line;
                if (n + 1 < sizeof(tok))
                    tok[n++] = *p;
                ++p;
            }
        } else {
            while ((*p >= '0' && *p <= '9') || (*p >= 'a' && *p <= 'z') || (*p >= 'A' && *p <= 'Z') ||
                   *p == '-' || *p == '.' || *p == '_' || *p == '+') {
                if (n + 1 < sizeof(tok))
                    tok[n++] = *p;
                ++p;
            }
            if (n == 0)
                return T_ERR; /* a character no token starts with */
        }
        tok[n] = 0;
        return T_STR;
    };
    auto fail = [&]() {
        if (err_line)
            *err_line = line;
        return (int)TBF_ERR_ARG;
    };
    int tk = next();
    while (tk != T_EOF) {
        int pgmnr;
        if (tk != T_STR || sscanf(tok, "%d", &pgmnr) != 1)
            return fail();
        if ((tk = next()) != '{')
            return fail();
        tk = next();
        while (tk != '}') {
            char sym[64];
            if (tk != T_STR)
                return fail();
            snprintf(sym, sizeof(sym), "%s", tok);
            if ((tk = next()) != '=')
                return fail();
            if ((tk = next()) != T_STR)
                return fail();
            if (tbf_programme_bind(t, pgmnr, sym, tok) != TBF_OK)
                return fail();
            tk = next();
            if (tk == ',')
                tk = next();
        }
        tk = next();
    }
    return TBF_OK;
}

unsigned tbf_sizeof_midi_map(void) { return (unsigned)sizeof(tbf_midi_map); }
unsigned tbf_sizeof_midi_state(void) { return (unsigned)sizeof(tbf_midi_state); }
unsigned tbf_sizeof_programme_table(void) { return (unsigned)sizeof(tbf_programme_table); }

} /* extern "C" */
