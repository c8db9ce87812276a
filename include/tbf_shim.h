/* tbf_shim.h -- the reference's per-instance engine API (struct b_tonegen / struct b_whirl and
 * the functions the LV2 / CLAP / JACK hosts call on them) backed by the CUDA path.
 *
 * libtbf_shim.so exports the reference's own symbol names with the reference's signatures
 * (src/tonegen.h:565-597, src/vibrato.h:77-79, src/whirl.h:223-272), so a host's synthSound()
 * (src/clap.cpp:244-270, b_synth/lv2.cpp:212-239) links against it instead of tonegen.o /
 * vibrato.o / whirl.o without source changes to the audio callback.  The structs are opaque:
 * each one owns one organ of a process-wide one-GPU tbf_ctx (include/tbf_gpu.h); the tone
 * generator half and the Leslie half are separate objects, as in struct b_instance
 * (src/global_inst.h:32-40), so preamp / reverb stay on the CPU between them.
 *
 * One organ per call is latency-bound on a GPU (one launch chain per fragment): this library
 * is the drop-in demonstration of the boundary; tbf_gpu_render_batch is the product.
 *
 * Not provided (control plane, SURVEY 8 out of scope): MIDI controller registration (`m` is
 * ignored), the fset/iset filter setters (send TBF_EV_HORN_FILTER through the batch API
 * instead).  On any failure the audio calls write silence (hosts never poll a status).
 */
#ifndef TBF_SHIM_H
#define TBF_SHIM_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

struct b_tonegen; /* opaque here; src/tonegen.h:176-563 there */
struct b_whirl;   /* src/whirl.h:65-221 */

/* the hosts' handle bag, src/global_inst.h:32-40, same layout; this library reads `synth` */
typedef struct b_instance {
    struct b_reverb *reverb;
    struct b_whirl *whirl;
    struct b_tonegen *synth;
    struct b_programme *progs;
    void *midicfg;
    void *preamp;
    void *state;
} b_instance;

/* src/cfgParser.h:39-44, same layout */
typedef struct _configContext {
    const char *fname;
    int linenr;
    const char *name;
    const char *value;
} ConfigContext;

/* ---- tone generator + scanner: src/tonegen.h:565-597, src/vibrato.h:77-79 ---- */
struct b_tonegen *allocTonegen(void);                          /* src/tonegen.cpp:3780 */
int oscConfig(struct b_tonegen *t, ConfigContext *cfg);        /* src/tonegen.cpp:2173 */
int scannerConfig(void *t, ConfigContext *cfg);                /* src/vibrato.cpp:334  */
/* `m` (the MIDI configuration) is ignored; targetRatio may be NULL (src/tonegen.cpp:2905) */
void initToneGenerator(struct b_tonegen *t, void *m, double rate, double *targetRatio);
void initVibrato(void *t, void *m);                            /* src/vibrato.cpp:319  */
void freeToneGenerator(struct b_tonegen *t);                   /* src/tonegen.cpp:3079 */
void oscKeyOn(struct b_tonegen *t, short midiNote, short realKey);  /* :3134 */
void oscKeyOff(struct b_tonegen *t, short midiNote, short realKey); /* :3096 */
void setDrawBar(struct b_tonegen *t, int bus, unsigned int setting); /* :2738 */
/* :2763; inst = the host's struct b_instance*, as in the reference: only ->synth is read */
void setDrawBars(void *inst, unsigned int manual, unsigned int setting[]);
void setPercussionEnabled(struct b_tonegen *t, int isEnabled); /* :1678 */
void setPercussionVolume(struct b_tonegen *t, int isSoft);     /* :1744 */
void setPercussionFast(struct b_tonegen *t, int isFast);       /* :1733 */
void setPercussionFirst(struct b_tonegen *t, int isFirst);     /* :1760 */
void setVibratoUpper(struct b_tonegen *t, int isEnabled);      /* :1635 */
void setVibratoLower(struct b_tonegen *t, int isEnabled);      /* :1650 */
void setVibratoFromInt(void *t, int param);                    /* src/vibrato.cpp:118  */
/* lengthSamples is ignored like in the reference: always one 128-sample fragment (:3239) */
void oscGenerateFragment(struct b_tonegen *t, float *buf, size_t lengthSamples);

/* ---- Leslie: src/whirl.h:223-272 ---- */
struct b_whirl *allocWhirl(void);                              /* src/whirl.cpp:136  */
int whirlConfig(struct b_whirl *w, ConfigContext *cfg);        /* src/whirl.cpp:992  */
void initWhirl(struct b_whirl *w, void *m, double rate);       /* src/whirl.cpp:956  */
void freeWhirl(struct b_whirl *w);
void useRevOption(struct b_whirl *w, int n, int signals);      /* src/whirl.cpp:174  */
void setRevSelect(struct b_whirl *w, int n);                   /* src/whirl.cpp:226  */
/* bufferLengthSamples must be a multiple of 128 (the hosts pass 128) */
void whirlProc(struct b_whirl *w, const float *inbuffer, float *outbL, float *outbR,
               size_t bufferLengthSamples);                    /* src/whirl.cpp:1640 */
/* NULL outputs are skipped (src/whirl.cpp:1191, src/whirl.h:232-238) */
void whirlProc2(struct b_whirl *w, const float *inbuffer, float *outL, float *outR, float *outHL,
                float *outHR, float *outDL, float *outDR, size_t bufferLengthSamples);
void whirlProc3(struct b_whirl *w, const float *inbuffer, float *outL, float *outR, float *tmpL,
                float *tmpR, size_t bufferLengthSamples);      /* src/whirl.cpp:1653 */

/* ---- what the reference takes from process globals ---- */
/* the 128 note frequencies the reference polls from MTS-ESP (src/tuning.cpp:25-34); call
 * before initToneGenerator; default 12-TET at A4 = 440 Hz */
void tbf_shim_set_tuning(struct b_tonegen *t, const double freqs[128]);
/* the srand() value in force when the reference's initToneGenerator runs (dither and click
 * envelopes consume rand(): src/tonegen.cpp:1450, :2577-2660); default 1 */
void tbf_shim_set_seed(struct b_tonegen *t, unsigned seed);
/* 0 = fine; otherwise the tbf_status of the first failed call on any object, with text */
int tbf_shim_status(void);
const char *tbf_shim_last_error(void);
/* tears the process-wide context down (all objects must have been freed) */
void tbf_shim_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* TBF_SHIM_H */
