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
 *   tbf_midi_state  the one piece of engine state a controller reads back: b_whirl::revSelect
 *                   (rotary.speed-toggle, src/whirl.cpp:252-261)
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
} tbf_midi_map;

typedef struct tbf_midi_state {
    int32_t rev_select;     /* b_whirl::revSelect: 0 slow, 1 stop, 2 fast (whirl.h:252-254) */
    uint32_t n_ignored;     /* messages with no effect on the path (reverb.mix, overdrive.*,
                               unassigned controllers, pitch bend, bank select ...)        */
    uint32_t n_unsupported; /* messages that DO change the reference's hot path but have no
                               tbf_event: program change, whirl.horn.filter.*, brake and
                               acceleration controllers.  Parity with the reference holds
                               only while this stays 0                                    */
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

void tbf_midi_state_init(tbf_midi_state *s, int32_t init_rev_select);

/* parse_raw_midi_data() + process_midi_event(): one raw message (2 or 3 bytes, status byte
 * first) that arrives before `fragment` of `instance`.  Appends the engine calls it stands for
 * to out[0..cap) and returns their number (0 for messages without effect; CC 120 / 123, the
 * panic, needs TBF_MAX_KEYS), or TBF_ERR_ARG when cap is too small or an argument is NULL. */
int tbf_midi_compile(const tbf_midi_map *m, tbf_midi_state *s, uint32_t instance, uint32_t fragment,
                     const uint8_t *msg, size_t size, tbf_event *out, size_t cap);

/* ccFuncNames[] (src/midi.cpp:76-187) and getCCFunctionId()'s first-prefix-match rule */
int tbf_midi_function_count(void);
const char *tbf_midi_function_name(int id);
int tbf_midi_function_id(const char *name);

unsigned tbf_sizeof_midi_map(void);
unsigned tbf_sizeof_midi_state(void);

#ifdef __cplusplus
}
#endif
#endif
