/* Device-side data layout shared by the kernels and the C-ABI host code.
 *
 * Per-instance run-time state is one flat POD blob in HBM.  Field meanings follow
 * struct b_tonegen (reference src/tonegen.h:176-563), struct b_vibrato
 * (src/vibrato.h:42-68) and struct b_whirl (src/whirl.h:65-221).
 *
 * The path runs as four stage kernels per chunk of fragments (tbf_kernels.cu):
 *   T  tone generator + scanner      warp = instance, lane = 4 samples of a fragment
 *   R  gain ramps, horn biquads,     lane = instance (the recurrences of 32 organs
 *      rotor speed ramp + angles     advance in lockstep in one warp)
 *   W  Leslie taps + delay lines     warp = instance, lane = sample of a 32-sample block
 *   D  drum biquads + output mix     lane = instance, then lane = sample for the mix
 * Every scalar of the state is owned by exactly one stage (the sub-structs below), so
 * stages of different chunks may overlap on different streams.
 */
#ifndef TBF_DEVICE_H
#define TBF_DEVICE_H

#include <stdint.h>
#include <cstddef>
#include <cuda.h>
#include "tbf_gpu.h"

/* rendering flag bits, src/tonegen.cpp:124-130 */
#define ORF_MODIFIED 0x4u
#define ORF_ADDED 0x2u
#define ORF_REMOVED 0x1u
#define OR_ADD 0x6u
#define OR_REM 0x5u
/* routing bits, src/tonegen.cpp:133-140 */
#define RT_PERC 0x0Cu
#define RT_UPPRVIB 0x02u
#define RT_LOWRVIB 0x01u
#define RT_VIB 0x03u

#define TBF_REC_CAP 160 /* longest active list whose records persist across fragments and launches */
#define TBF_VRING 64 /* shared-memory scanner ring: host checks that every offset-table
                        entry keeps the writes inside it (tbf_api.cu); offsets are
                        (1 + A + A sin) with A <= 12 (scanner.modulation.*'s range,
                        src/vibrato.cpp:343-352): 25 slots ahead + a 32-sample block + 2 */

struct OscState {      /* struct _oscillator + the scalar part of AOTElement */
    uint32_t pos;      /* wave read position                                 */
    int32_t aclPos;    /* position in active list, -1 = not listed           */
    uint32_t rflags;
    int32_t refCount;
    float sumUpper, sumLower, sumPedal, sumPercn, sumSwell, sumScanr;
};

/* ---- stage T: tone generator and scanner ---- */
struct TgScalars {
    int32_t activeOscLEnd;
    uint32_t newRouting, oldRouting;
    int32_t percSendBus;
    uint32_t upperKeyCount;
    int32_t keyDownCount;
    float swellPedalGain;
    uint32_t drawBarChange;
    int32_t percEnabled, percTrigRestore, percIsSoft, percIsFast;
    float percEnvGainReset, percEnvGainDecay, percDrawbarGain;
    float outputGain, pz;
    float drawBarGain[TBF_NOF_BUSES];
    uint32_t vibStator, vibOutPos;
    int32_t vibTable, vibMixed;
    uint32_t errorFlags;
    int32_t tgTables, whTables; /* read-only after create; every stage reads them */
    uint32_t pad_;
    unsigned long long statFragments, statActiveWheels;
};

/* ---- stage R: percussion / key-compression ramps, horn filters, rotors ---- */
struct RampScalars {
    float percEnvGain, keyCompLevel;
    float iirz[2][2]; /* hafw hbfw : z0 z1 */
    float iir[2][5];  /* hafw hbfw : a1 a2 b0 b1 b2; the table set's until TBF_EV_HORN_FILTER */
    int32_t hornAcDc, drumAcDc, revSelect;
    uint32_t errorFlags;
    double hornAngle, drumAngle, hornIncr, drumIncr, hornTarget, drumTarget;
    double brakePos[2]; /* hnBrakePos drBrakePos; TBF_EV_WHIRL_BRAKE                     */
    double oml[4];      /* DevWhirlTables::one_minus_l until TBF_EV_WHIRL_RAMP           */
};

/* ---- stage W: Leslie reflections, taps, delay lines ---- */
struct WhirlScalars {
    uint32_t outpos;      /* the horn half's copy of the delay lines' read position */
    uint32_t errorFlags;
    float z[4];
    float adx[3][8];
    int32_t adi[3];
    uint32_t outpos_drum; /* the drum half's copy: the two halves of stage W run as independent
                             warps, each reads and advances its own                            */
};

/* ---- stage D: drum filters ---- */
struct DrumScalars {
    float iirz[2][2]; /* drfL drfR : z0 z1 */
};

struct InstState {
    TgScalars tg;
    RampScalars rp;
    WhirlScalars wl;
    DrumScalars dr;
    int32_t activeOscList[TBF_NOF_WHEELS + 1];
    uint8_t activeKeys[TBF_MAX_KEYS];
    OscState osc[TBF_NOF_WHEELS + 1];
    float busLevel[(TBF_NOF_WHEELS + 1) * TBF_NOF_BUSES]; /* aot[w].busLevel[b] */
    float vibBuffer[TBF_VIB_BUF];
    float wbuf[4][TBF_WHIRL_BUF]; /* HL HR DL DR */
    /* Stage T's records of the active list (gains, wave length | position of the fragment last
     * rendered, wheel), kept from one launch to the next: they are derived state, valid only while
     * rec_valid is set (stage T clears it whenever its last fragment left no reusable records; a
     * fresh or re-initialised instance has it clear).  With them the first fragment of a chunk is as
     * cheap as any other quiet one instead of a full active-list pass over state in HBM. */
    alignas(16) float recG[TBF_REC_CAP][4];
    uint32_t recL[TBF_REC_CAP];
    uint16_t recW[TBF_REC_CAP];
    int32_t rec_valid, rec_count;
};
static_assert(offsetof(InstState, recG) % 16 == 0 && sizeof(InstState) % 16 == 0, "recG is read as float4");

/* Floats per rotated copy of a wave of `len` samples: copy c holds wave[(i + c) % len] for
 * i < len + 256, padded to whole float4s, so the reads of TWO consecutive fragments from any
 * position are one contiguous, 16-byte aligned kilobyte.  Copies 0..3 of a wheel lie back to back. */
static inline __host__ __device__ uint32_t tbf_wave_copy_len(uint32_t len)
{
    return (len + 2u * TBF_FRAGMENT + 3u) & ~3u;
}

/* tbf_tonegen_tables with device pointers */
struct DevTonegenTables {
    const float *wave4; /* four rotated, wrap-extended copies per wave, see tbf_wave_copy_len() */
    const uint16_t *contrib_wheel;
    const uint16_t *contrib_bus;
    const float *contrib_level;
    /* per key: runs of contributions to one wheel, (wheel, entries) in list order; lets a
     * warp apply a key event with lane = contribution / lane = wheel */
    const uint16_t *run_wheel;
    const uint16_t *run_count;
    uint32_t run_start[TBF_MAX_KEYS + 1];
    int32_t coop_events; /* 1: every key's list is wheel-sorted with unique (wheel, bus) pairs */
    uint32_t wave4_offset[TBF_NOF_WHEELS + 1]; /* floats, multiple of 4: copy 0 of the wheel */
    uint32_t wave_length[TBF_NOF_WHEELS + 1];
    uint32_t contrib_start[TBF_MAX_KEYS + 1];
    float env[2 * 8][TBF_FRAGMENT]; /* rows 0..7 attack, 8..15 release (row 8 of the
                                       reference's 9 is never used: index is i & 7) */
    float key_comp_table[128];
    float draw_bar_level[TBF_NOF_BUSES][TBF_NOF_DRAWBARS];
    float perc_env_gain_reset_norm, perc_env_gain_reset_soft, perc_env_scaling;
    float perc_decay_fast_norm, perc_decay_fast_soft, perc_decay_slow_norm, perc_decay_slow_soft;
    float perc_drawbar_normal_gain, perc_drawbar_soft_gain;
    int32_t perc_send_bus_a, perc_send_bus_b, perc_trigger_bus;
    float output_level_trim;
    uint32_t vib_offset_table[3][TBF_VIB_TBL];
    uint32_t vib_stator_increment;
};

struct DevWhirlTables {
    /* displacement tables as pairs {d[k], d[(k + 1) & mask]} (one 8-byte load per tap) */
    const float2 *hn_fwd_displ, *hn_bwd_displ, *dr_fwd_displ, *dr_bwd_displ;
    const float *bfw, *bbw; /* FIR banks, five planes of 16384: coefficient c of position k at [c][k] */
    double sample_rate;
    double fw_ang, bw_ang;          /* micAngle*.25, 1 + micAngle*-.25  (whirl.cpp:1379-1380) */
    double one_minus_l[4];          /* 1 - exp(-1/(SR/128*tau)), tau = hornAcc hornDec drumAcc
                                       drumDec (whirl.cpp:1255, :1306), glibc exp on the host */
    /* the same for every value a ramp controller can send: tau = (float)(.01 + v / 80.0)
     * for the horn's two, (float)(.01 + v / 14.0) for the drum's (whirl.cpp:895-917)     */
    double ramp_one_minus_l[4][128];
    double rev_horn_target[9], rev_drum_target[9];
    double hn_brake_pos, dr_brake_pos;
    int32_t horn_phase[6], drum_phase[6];
    float horn_spacing[6], drum_spacing[6];
    float iir[4][5];                /* hafw hbfw drfL drfR : a1 a2 b0 b1 b2 */
    int32_t revselects[3];
    float horn_level, leakage;
    float mic_hll, mic_hlr, mic_hrl, mic_hrr, mic_dll, mic_dlr, mic_drl, mic_drr;
    int32_t bypass;
    /* delay-line geometry derived from the tables on the host (tbf_api.cu):
     * every tap writes `delay` slots ahead of the reader with
     * tap_dmin <= delay <= tap_dmax (secondary slot included) */
    int32_t ring_size;      /* 512, 1024 or 2048: shared-memory ring per delay line       */
    int32_t gather_ok;      /* 1: the three taps of every line have disjoint delay ranges,
                               so the ordered scatter can be evaluated in parallel passes  */
    int32_t window;         /* slots ahead of the reader that can be non-zero after a
                               32-sample block has been scattered                         */
    int32_t small_index;    /* 1: every phase is in [0, 2^21) and every tap delay non-negative, so
                               with the rotor angle in [+0, fast_max_angle) the table index
                               h1 = angle * 16384 + phase and the delay-line position t lie in
                               [+0, 2^22): stage W reads floor / round / the integer conversions
                               off the float's mantissa instead of converting (tbf_kernels.cu)  */
    double fast_max_angle;  /* 1.0: horn angle + microphone offset stays below 2^7            */
    int32_t taps_certified; /* 1: the host has bounded every tap's delay change per sample by
                               0.49 for every rotor speed this set can reach (tbf_api.cu), so
                               the per-block slot checks of the ordered scatter cannot fire and
                               stage W skips them                                               */
    int32_t pad_cert_;
    int32_t tap_order[4][3];/* per line HL HR DL DR: tap ids (0..11), farthest delay first */
    int32_t tap_dmin[12], tap_dmax[12];
};

/* what stage T hands to stage R for every (slot, fragment) */
struct __align__(16) FragCtl {
    float key_comp;  /* keyCompTable[keyDownCount]                           */
    float gain;      /* swellPedalGain * percDrawbarGain, or swellPedalGain   */
    float decay;     /* percEnvGainDecay                                     */
    float reset;     /* percEnvGainReset                                     */
    uint32_t flags;  /* FC_*                                                 */
    uint32_t pad_[3];
};
#define FC_PERC_ON 0x1u     /* RT_PERC routed: percussion term + decay         */
#define FC_PERC_RESET 0x2u  /* no upper key down at the end of the fragment    */

struct RenderParams {
    InstState *states;
    const uint32_t *ids;       /* NULL = identity: instance = id_base + slot */
    uint32_t id_base;
    const DevTonegenTables *tg;
    const DevWhirlTables *wh;
    const tbf_event *events;
    const uint64_t *ev_start;  /* [n_instances + 1] */
    uint64_t *ev_cursor_t;     /* [n_instances]: next unapplied event of stage T, over chunks */
    uint64_t *ev_cursor_r;     /* same for stage R (it applies the rotor-speed events)        */
    uint64_t *ev_cursor_rot;   /* the rotor kernel's own cursor when it runs ahead of stage R's audio chain (else NULL) */
    float *out_l, *out_r;
    float *stems[4];           /* TBF_RENDER_STEMS: outHL outHR outDL outDR, or NULL */
    unsigned long long *checksums; /* [n_instances] or NULL */
    uint64_t out_stride;
    uint32_t n_instances, n_fragments, flags;
    uint32_t frag_base;        /* batch fragment index of this launch's fragment 0 */
    /* stage-to-stage streams of this chunk, [slot][pitch] floats (doubles for angles) */
    FragCtl *ctl;              /* [slot][n_fragments]                                   */
    float *u, *php;            /* T -> R: swell (+ scanner) sum, percussion high-pass   */
    float *dx0, *hx0;          /* R -> W: whirl input, horn-filter output                */
    double *hang, *dang;       /* R -> W: rotor angles                                  */
    float *line[4];            /* W -> D: delay-line reads HL HR DL DR                  */
    uint64_t pitch;            /* elements per slot row of every stream                 */
    float one;                 /* 1.0f, as a run-time value (tbf_kernels.cu: add2_via_fma) */
    uint32_t tg_sets;          /* tone-generator table sets the batch's organs share (0: unknown) */
    /* 2-D TMA tiles (stages R and D): the launch's first slot as a row of the tensor maps, per plane
     * u php dx0 hx0 line[0..3] (float map) and hang dang (double map); tile_maps = host pointer to the
     * maps, NULL when the scratch layout has none (then the kernels move tiles with cp.async) */
    uint32_t tile_row[10];
    const struct TileMaps *tile_maps;
};
/* tensor maps over the stage-stream scratch, rows = every plane's slots back to back */
struct TileMaps {
    CUtensorMap m32; /* float planes: dim0 = pitch, dim1 = rows */
    CUtensorMap m64; /* double planes as pairs of floats: dim0 = 2 * pitch */
};

/* what a fresh instance starts from (tbf_init_instances_kernel); everything else is zero */
struct InstInit {
    TgScalars tg;
    RampScalars rp;
};
/* kernel-side view of tbf_wave_synth (include/tbf_tables.h) */
struct WaveSynthParams {
    double rate;
    double harmonics[TBF_MAX_PARTIALS];
};

size_t tbf_stage_t_smem(void);
size_t tbf_stage_w_smem(int ring_size);
int tbf_launch_stage_t(const RenderParams &p, void *stream);
int tbf_launch_stage_r(const RenderParams &p, void *stream);
bool tbf_stage_r_rotors_apart(const RenderParams &p);
int tbf_launch_stage_rot(const RenderParams &p, void *stream); /* before stage R waits for stage T */
int tbf_launch_stage_w(const RenderParams &p, int ring_size, void *stream);
int tbf_launch_stage_d(const RenderParams &p, void *stream);
int tbf_launch_wave_store(float *wave4, const float *raw, const uint32_t *meta, void *stream);
int tbf_launch_replicate(InstState *states, uint32_t first, uint32_t count, const InstState *dev_tmpl,
                         void *stream);
int tbf_launch_init_instances(InstState *states, uint32_t first, uint32_t count, const InstInit &init,
                              void *stream);
int tbf_launch_wave_synth(float *raw, const uint32_t *meta, const double *wheel_f1, const double *wheel_att,
                          const WaveSynthParams &sp, const uint32_t *dither_bits, uint32_t *flag_count,
                          void *flags, uint32_t flag_cap, void *stream);

#endif
