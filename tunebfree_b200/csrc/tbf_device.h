/* Device-side data layout shared by the kernels and the C-ABI host code.
 *
 * Per-instance run-time state is one flat POD blob in HBM (array of structs:
 * one CTA owns one instance, so its threads read consecutive words of one
 * blob).  Field meanings follow struct b_tonegen (reference src/tonegen.h:176-563),
 * struct b_vibrato (src/vibrato.h:42-68) and struct b_whirl (src/whirl.h:65-221).
 */
#ifndef TBF_DEVICE_H
#define TBF_DEVICE_H

#include <stdint.h>
#include "tbf_gpu.h"

#define TBF_THREADS 128

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

struct OscState {      /* struct _oscillator + the scalar part of AOTElement */
    uint32_t pos;      /* wave read position                                 */
    int32_t aclPos;    /* position in active list, -1 = not listed           */
    uint32_t rflags;
    int32_t refCount;
    float sumUpper, sumLower, sumPedal, sumPercn, sumSwell, sumScanr;
};

/* Every scalar of the run-time state: copied to shared memory when a launch
 * starts and back when it ends, so the serial control code never waits on HBM. */
struct InstScalars {
    /* ---- tone generator ---- */
    int32_t activeOscLEnd;
    uint32_t newRouting, oldRouting;
    int32_t percSendBus;
    uint32_t upperKeyCount;
    int32_t keyDownCount;
    float swellPedalGain;
    uint32_t drawBarChange;
    int32_t percEnabled, percTrigRestore, percIsSoft, percIsFast;
    float percEnvGain, percEnvGainReset, percEnvGainDecay, percDrawbarGain;
    float outputGain, pz, keyCompLevel;
    float drawBarGain[TBF_NOF_BUSES];
    /* ---- scanner vibrato ---- */
    uint32_t vibStator, vibOutPos;
    int32_t vibTable, vibMixed;
    /* ---- whirl ---- */
    int32_t hornAcDc, drumAcDc, revSelect;
    uint32_t outpos;
    double hornAngle, drumAngle, hornIncr, drumIncr, hornTarget, drumTarget;
    float z[4];
    float iirz[4][2]; /* hafw hbfw drfL drfR : z0 z1 */
    float adx[3][8];
    int32_t adi[3];
    /* ---- bookkeeping ---- */
    uint32_t errorFlags;
    int32_t tgTables, whTables;
    uint32_t pad_;
    unsigned long long statFragments, statActiveWheels; /* fragments rendered, sum of active-list lengths */
};

struct InstState {
    InstScalars sc;
    int32_t activeOscList[TBF_NOF_WHEELS + 1];
    uint8_t activeKeys[TBF_MAX_KEYS];
    OscState osc[TBF_NOF_WHEELS + 1];
    float busLevel[(TBF_NOF_WHEELS + 1) * TBF_NOF_BUSES]; /* aot[w].busLevel[b] */
    float vibBuffer[TBF_VIB_BUF];
    float wbuf[4][TBF_WHIRL_BUF]; /* HL HR DL DR */
};

/* tbf_tonegen_tables with device pointers */
struct DevTonegenTables {
    const float *wave_data;
    const uint16_t *contrib_wheel;
    const uint16_t *contrib_bus;
    const float *contrib_level;
    uint32_t wave_offset[TBF_NOF_WHEELS + 1];
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
    const float *hn_fwd_displ, *hn_bwd_displ, *dr_fwd_displ, *dr_bwd_displ;
    const float *bfw, *bbw;
    double sample_rate;
    double fw_ang, bw_ang;          /* micAngle*.25, 1 + micAngle*-.25  (whirl.cpp:1379-1380) */
    double one_minus_l[4];          /* 1 - exp(-1/(SR/128*tau)), tau = hornAcc hornDec drumAcc
                                       drumDec (whirl.cpp:1255, :1306), glibc exp on the host */
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
    int32_t ring_size;      /* 512 or 2048: shared-memory ring per delay line          */
    int32_t gather_ok;      /* 1: the three taps of every line have disjoint delay ranges,
                               so the ordered scatter can be evaluated as a parallel gather */
    int32_t window;         /* slots ahead of the reader that can be non-zero, + 128    */
    int32_t tap_order[4][3];/* per line HL HR DL DR: tap ids (0..11), farthest delay first */
    int32_t tap_dmin[12], tap_dmax[12];
    int32_t line_dmin[4];   /* per line: slots nearer than this cannot be written in a fragment */
};

struct RenderParams {
    InstState *states;
    const uint32_t *ids;       /* NULL = identity */
    const DevTonegenTables *tg;
    const DevWhirlTables *wh;
    const tbf_event *events;
    const uint64_t *ev_start;  /* [n_instances + 1] */
    uint64_t *ev_cursor;       /* [n_instances]: next unapplied event, carried over chunks */
    float *out_l, *out_r;
    unsigned long long *checksums; /* [n_instances] or NULL */
    uint64_t out_stride;
    uint32_t n_instances, n_fragments, flags;
    uint32_t frag_base;        /* batch fragment index of this launch's fragment 0 */
};

size_t tbf_render_smem_bytes(int ring_size);
int tbf_launch_render(const RenderParams &p, int ring_size, void *stream);
int tbf_launch_replicate(InstState *states, uint32_t first, uint32_t count, const InstState *dev_tmpl,
                         void *stream);

#endif
