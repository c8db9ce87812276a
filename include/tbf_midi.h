/* tbf_midi.h -- raw MIDI -> (fragment, event) compiler for the batched renderer.
 *
 * The reference hosts hand raw MIDI messages to parse_raw_midi_data()
 * (src/midi.cpp:1261-1302) between two 128-sample fragments; process_midi_event()
 * (src/midi.cpp:1095-1255) maps notes through the per-channel key tables and controllers
 * through the controller -> function tables and calls the engine (oscKeyOn/oscKeyOff,
 * setDrawBar, setPercussion*, setVibrato*, useRevOption ...) directly.  The batched renderer
 * takes a sorted event list instead (tbf_event, tbf_gpu.h); this module is the host-side
 * translation of one into the other, per organ instance:
 *
 *   tbf_midi_map    what resetMidiCfg() + midiPrimeControllerMapping() + midiConfig() +
 *                   initMidiTables() + the modules' useMIDIControlFunction() calls leave in
 *                   struct b_midicfg (src/midi.cpp:203-268): receive channels, transposes and
 *                   splits, function -> controller tables, invert flags; and, derived by
 *                   tbf_midi_map_finish(), note -> key and (channel, controller) -> function
 *   tbf_midi_state  the engine state a controller reads back: b_whirl::revSelect
 *                   (rotary.speed-toggle, src/whirl.cpp:252-261), the vibrato routing, the key
 *                   mapping a programme may move, and the horn filters' type / Hz / Q / gain
 *                   (each whirl.horn.filter.* controller changes one and recomputes the biquad)
 *
 * Plain C ABI, no allocation, no globals; one map may serve any number of instances.
 */
#ifndef TBF_MIDI_H
#define TBF_MIDI_H

#include <stddef.h>
#include <stdint.h>

#include "tbf_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

#define TBF_MIDI_MAX_FN 80   /* >= number of controllable functions (ccFuncNames, src/midi.cpp:76-187) */
#define TBF_MIDI_UNMAPPED 255
#define TBF_MIDI_FLAG_INV 1  /* MFLAG_INV, src/midi.h: value -> 127 - value */

/* ---- programmes (src/program.h:36-108, src/program.cpp, src/pgmParser.cpp) ---- */
#define TBF_MAXPROGS 129 /* MAXPROGS */
/* tbf_programme::flags, the FL_* bits of src/program.h:66-98 */
#define TBF_PGM_INUSE 0x0001u
#define TBF_PGM_DRAWBR 0x0002u
#define TBF_PGM_SCANNR 0x0100u
#define TBF_PGM_PRCENA 0x0200u
#define TBF_PGM_PRCVOL 0x0400u
#define TBF_PGM_PRCSPD 0x0800u
#define TBF_PGM_PRCHRM 0x1000u
#define TBF_PGM_OVRSEL 0x2000u
#define TBF_PGM_ROTENA 0x4000u
#define TBF_PGM_ROTSPS 0x8000u
#define TBF_PGM_RVBMIX 0x00010000u
#define TBF_PGM_DRWRND 0x00020000u
#define TBF_PGM_KSPLTL 0x00040000u
#define TBF_PGM_LOWDRW 0x00080000u
#define TBF_PGM_PDLDRW 0x00100000u
#define TBF_PGM_KSPLTP 0x00200000u
#define TBF_PGM_TRA_PD 0x00400000u
#define TBF_PGM_TRA_LM 0x00800000u
#define TBF_PGM_TRA_UM 0x01000000u
#define TBF_PGM_TRANSP 0x02000000u
#define TBF_PGM_TRCH_A 0x04000000u
#define TBF_PGM_TRCH_B 0x08000000u
#define TBF_PGM_TRCH_C 0x10000000u
#define TBF_PGM_VCRUPR 0x20000000u
#define TBF_PGM_VCRLWR 0x40000000u

typedef struct tbf_programme {
    uint32_t flags;
    uint8_t drawbars[3][9];   /* upper, lower, pedals: 0..8                                 */
    uint8_t perc_enabled, perc_soft, perc_fast, perc_second, overdrive;
    int16_t scanner;          /* low byte VIB1..3 = 1..3, CHO1..3 = 0x81..0x83; 0x200 upper,
                                 0x100 lower routing (src/vibrato.h:30-36)                  */
    int16_t rotary_speed;     /* 0 slow, 1 stop, 2 fast                                     */
    int16_t split_lower, split_pedals;
    int16_t transpose[7];     /* global, channel A, B, C, split upper, split pedals, split
                                 lower -- own order, see tbf_programme_bind()               */
    float reverb_mix;
    char name[22];
} tbf_programme;

typedef struct tbf_programme_table {
    int32_t controller_offset; /* MIDIControllerPgmOffset, 1 unless pgm.controller.offset=0 */
    int32_t previous;          /* previousPgmNr of bindToProgram                            */
    tbf_programme prog[TBF_MAXPROGS];
} tbf_programme_table;

/* An EMPTY table (the reference starts from its compiled-in copy of pgm/default.pgm, which is
 * not part of this library: load the programme text you want served). */
void tbf_programmes_init(tbf_programme_table *t);
/* One property of one programme, bindToProgram() (src/program.cpp:308-702): same keywords
 * and value spellings.  0 or TBF_ERR_ARG. */
int tbf_programme_bind(tbf_programme_table *t, int pgmnr, const char *sym, const char *val);
/* A programme file in the syntax of src/pgmParser.cpp:65-75:
 *   <number> '{' <property> '=' <value> [[,] ...] '}' ..., '#' comments, "quoted strings".
 * Returns 0, or TBF_ERR_ARG with the 1-based line of the first error in *err_line. */
int tbf_programmes_load(tbf_programme_table *t, const char *text, int *err_line);

typedef struct tbf_midi_map {
    /* -- configuration (struct b_midicfg, src/midi.cpp:203-268) -- */
    uint8_t rcv_ch[3];      /* rcvChA, rcvChB, rcvChC: upper, lower, pedals; 0..15        */
    uint8_t pad_;
    int32_t transpose;      /* all channels                                               */
    int32_t nsh_a, nsh_a_u, nsh_a_pl, nsh_a_ul, nsh_b, nsh_c;
    int32_t split_a_pl, split_a_ul;
    uint8_t cc_use[3][TBF_MIDI_MAX_FN]; /* ctrlUseA/B/C: function id -> controller, 255 = none */
    uint8_t cc_flags[16][128];          /* ctrlflg[channel][controller]                    */
    int32_t revselects[3];              /* b_whirl::revselects (tbf_whirl_tables)          */
    /* -- derived by tbf_midi_map_finish() -- */
    int16_t key_table[16][128];         /* keyTable[channel][note] -> engine key, -1 = none */
    int16_t cc_fn[16][128];             /* ctrlvec[channel][controller] -> function id, -1  */
    int8_t ch_set[16];                  /* channel -> 0 upper, 1 lower, 2 pedals, -1 none    */
    /* programme changes (installProgram, src/program.cpp:735-921) are served from here;
     * NULL: a programme change is counted as unsupported */
    const tbf_programme_table *programmes;
} tbf_midi_map;

typedef struct tbf_midi_state {
    int32_t rev_select;     /* b_whirl::revSelect: 0 slow, 1 stop, 2 fast (whirl.h:252-254) */
    int32_t vib_routing;    /* getVibratoRouting(): 1 lower, 2 upper (tonegen.cpp:1662)     */
    /* the instance's key mapping: starts as the map's, programmes may change it
     * (setKeyboardSplitMulti / setKeyboardTranspose*, src/midi.cpp:728-780) */
    int32_t transpose, nsh_a, nsh_a_u, nsh_a_pl, nsh_a_ul, nsh_b, nsh_c, split_a_pl, split_a_ul;
    /* haT haF haQ haG / hbT hbF hbQ hbG (src/whirl.h:173-180, floats there) and the rate
     * setIIRFilter() is called with (src/whirl.cpp:674-684) */
    float horn_filter[2][4];
    double sample_rate;
    uint32_t n_ignored;     /* messages with no effect on the path (reverb.mix, overdrive.*,
                               unassigned controllers, pitch bend, bank select ...)        */
    uint32_t n_unsupported; /* messages that DO change the reference's hot path but have no
                               tbf_event: programmes with random drawbars (process-global
                               rand()), programme changes without a table.  Parity with the
                               reference holds only while this stays 0                     */
    uint32_t n_unmapped;    /* notes whose key is TBF_MAX_KEYS or above after transposition:
                               no event, as oscKeyOn / oscKeyOff return early
                               (src/tonegen.cpp:3098, :3136)                               */
} tbf_midi_state;

/* resetMidiCfg() + clearControllerMapping() + midiPrimeControllerMapping()
 * (src/midi.cpp:270-296, :591-623, :839-900): the compiled-in defaults */
void tbf_midi_map_default(tbf_midi_map *m);

/* One `key=value` configuration line, midiConfig() (src/midi.cpp:906-1034): midi.*.channel,
 * midi.transpose, midi.{upper,lower,pedals}.transpose[.split], midi.controller.reset,
 * midi.controller.{upper,lower,pedals}.<cc>=<function>[-] | unmap.
 * Returns 1 if the line was consumed, 0 if it is not a midi.* key, TBF_ERR_ARG on a value the
 * reference rejects. */
int tbf_midi_map_apply_cfg(tbf_midi_map *m, const char *key, const char *value);

/* setKeyboardSplitMulti() (src/midi.cpp:728-744), flags as there: 1 split_a_pl, 2 split_a_ul,
 * 4 nsh_a_pl, 8 nsh_a_ul, 16 nsh_a_u */
void tbf_midi_map_split(tbf_midi_map *m, int flags, int split_a_pl, int split_a_ul, int nsh_a_pl,
                        int nsh_a_ul, int nsh_a_u);

/* initMidiTables() (src/midi.cpp:1037-1044: key tables A, B, C and the per-channel status
 * table) + the registration pass of the engine's init functions (useMIDIControlFunction,
 * src/midi.cpp:416-446, called in the order initToneGenerator, initVibrato, initPreamp,
 * initReverb, initWhirl: a later function takes a controller an earlier one held).
 * revselects = tbf_whirl_tables::revselects, NULL = the defaults {4, 0, 8}. */
int tbf_midi_map_finish(tbf_midi_map *m, const int32_t revselects[3]);

/* What a fresh instance of table set `w` starts from: init_rev_select, the horn filter
 * parameters and the sample rate (w = NULL: whirl.cpp's compiled-in defaults at 48 kHz);
 * the vibrato routing starts at 0, the key mapping at the map's */
void tbf_midi_state_init(tbf_midi_state *s, const tbf_midi_map *m, const tbf_whirl_tables *w);
/* engine key of (channel, note) under the instance's current mapping, -1 = none */
int tbf_midi_key(const tbf_midi_map *m, const tbf_midi_state *s, int channel, int note);

/* parse_raw_midi_data() + process_midi_event(): one raw message (2 or 3 bytes, status byte
 * first) that arrives before `fragment` of `instance`.  Appends the engine calls it stands for
 * to out[0..cap) and returns their number (0 for messages without effect; CC 120 / 123, the
 * panic, needs TBF_MAX_KEYS; a programme change up to 40; a horn filter controller 5: the
 * biquad's coefficients, computed here by tbf_iir_coefficients()), or TBF_ERR_ARG when cap is too
 * small or an argument is NULL. */
int tbf_midi_compile(const tbf_midi_map *m, tbf_midi_state *s, uint32_t instance, uint32_t fragment,
                     const uint8_t *msg, size_t size, tbf_event *out, size_t cap);

/* ccFuncNames[] (src/midi.cpp:76-187) and getCCFunctionId()'s first-prefix-match rule */
int tbf_midi_function_count(void);
const char *tbf_midi_function_name(int id);
int tbf_midi_function_id(const char *name);

unsigned tbf_sizeof_midi_map(void);
unsigned tbf_sizeof_midi_state(void);
unsigned tbf_sizeof_programme_table(void);

#ifdef __cplusplus
}
#endif
#endif
