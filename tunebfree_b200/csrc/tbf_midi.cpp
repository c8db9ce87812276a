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
    K_UNSUPPORTED,  /* registered, changes the path, no tbf_event                        */
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
        add_fn(n, K_UNSUPPORTED, 0, r_whirl + 3 + i);
    }
    static const char *df[4] = {"type", "hz", "q", "gain"};
    for (int i = 0; i < 4; ++i) {
        char n[32];
        snprintf(n, sizeof(n), "whirl.drum.filter.%s", df[i]);
        add_fn(n, K_UNSUPPORTED, 0, 0); /* named, but no module registers them */
    }
    add_fn("whirl.horn.brakepos", K_UNSUPPORTED, 0, r_whirl + 11);
    add_fn("whirl.drum.brakepos", K_UNSUPPORTED, 0, r_whirl + 12);
    add_fn("whirl.horn.acceleration", K_UNSUPPORTED, 0, r_whirl + 13);
    add_fn("whirl.horn.deceleration", K_UNSUPPORTED, 0, r_whirl + 14);
    add_fn("whirl.drum.acceleration", K_UNSUPPORTED, 0, r_whirl + 15);
    add_fn("whirl.drum.deceleration", K_UNSUPPORTED, 0, r_whirl + 16);
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

/* revSelect after useRevOption(w, n, 2) (whirl.cpp:205-218) */
inline int rev_select_of_option(int n)
{
    const int hr = (n / 3) % 3;
    return hr == 2 ? 2 /* WHIRL_FAST */ : (hr == 1 ? 0 /* WHIRL_SLOW */ : 1 /* WHIRL_STOP */);
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
    int16_t kt[3][128];
    for (int i = 0; i < 128; ++i) {
        int off;
        if (i < m->split_a_pl)
            off = m->transpose + m->nsh_a_pl + 2 * 128;
        else if (i < m->split_a_ul)
            off = m->transpose + m->nsh_a_ul + 128;
        else if (m->split_a_ul)
            off = m->transpose + m->nsh_a_u;
        else
            off = m->transpose + m->nsh_a;
        kt[0][i] = (int16_t)(i + off);
        kt[1][i] = (int16_t)(128 + i + m->transpose + m->nsh_b);
        kt[2][i] = (int16_t)(2 * 128 + i + m->transpose + m->nsh_c);
    }
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
    for (int s = 0; s < 3; ++s) {
        const int ch = m->rcv_ch[s] & 15;
        memcpy(m->key_table[ch], kt[s], sizeof(kt[s]));
        memcpy(m->cc_fn[ch], vec[s], sizeof(vec[s]));
    }
    return TBF_OK;
}

void tbf_midi_state_init(tbf_midi_state *s, int32_t init_rev_select)
{
    if (!s)
        return;
    memset(s, 0, sizeof(*s));
    s->rev_select = init_rev_select % 3;
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
    } while (0)

    if (status == 0x80 || status == 0x90) {
        const int note = msg[1] & 0x7F, vel = msg[2] & 0x7F;
        const int key = m->key_table[ch][note];
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
    if (status == 0xC0) { /* installProgram(): programme tables are outside this module */
        s->n_unsupported++;
        return 0;
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

unsigned tbf_sizeof_midi_map(void) { return (unsigned)sizeof(tbf_midi_map); }
unsigned tbf_sizeof_midi_state(void) { return (unsigned)sizeof(tbf_midi_state); }

} /* extern "C" */
