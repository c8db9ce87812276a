/* tbf_tables.h -- host-side table builder of the B200 renderer (C-ABI).
 *
 * Turns a resolved instance specification (128 MTS-ESP note frequencies,
 * osc.* / scanner.* / whirl.* parameters, rand() seed, sample rate) into the
 * immutable tables of include/tbf_gpu.h.  It is a from-scratch restatement of
 * the reference's init path:
 *   getFrequencies / extendFrequencies      src/tuning.cpp:48-147
 *   initToneGenerator and what it calls     src/tonegen.cpp:2905-3066
 *     applyDefaultConfiguration :941-1023, applyManualDefaults :707-802,
 *     compilePlayMatrix :1122-1213, initOscillators :1470-1630, fitWave :1335-1369,
 *     writeSamples :1402-1457, initKeyCompTable :1939-1965, initEnvelopes :2562-2728
 *   init_vibrato                            src/vibrato.cpp:224-317
 *   initWhirl                               src/whirl.cpp:338-662, :956-990
 *   eqCompute                               src/eqcomp.cpp:109-203
 * Init is cold code: it runs on the host with the same libm the reference uses
 *  and its own re-implementation of glibc's TYPE_3 rand().
 */
#ifndef TBF_TABLES_H
#define TBF_TABLES_H

#include "tbf_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

#define TBF_MAX_EXTRA_HARMONICS 64

/* ENV_* models, src/tonegen.h:43-47 */
#define TBF_ENV_CLICK 0
#define TBF_ENV_COSINE 1
#define TBF_ENV_LINEAR 2
#define TBF_ENV_SHELF 3
/* osc.eq.macro, src/tonegen.cpp:143-145 */
#define TBF_EQ_SPLINE 0
#define TBF_EQ_PEAK24 1
#define TBF_EQ_PEAK46 2

typedef struct tbf_tonegen_spec {
    double sample_rate;
    double freqs[128];       /* MTS_NoteToFrequency(note 0..127), src/tuning.cpp:25 */
    double target_ratio[9];  /* drawbar harmonic ratios, src/tonegen.cpp:2959      */
    uint32_t seed;           /* srand(seed) right before initToneGenerator          */
    /* osc.* (defaults: initValues src/tonegen.cpp:196-300) */
    double precision;        /* osc.x-precision                                     */
    int32_t eq_macro;
    double eq_p1y, eq_r1y, eq_p4y, eq_r4y;
    double compartment_crosstalk, transformer_crosstalk;
    double terminalstrip_crosstalk, wiring_crosstalk;
    double contribution_floor, contribution_min;
    int32_t n_extra_harmonics; /* osc.harmonic.<n>=v lines, in file order: they ADD  */
    int32_t extra_harmonic_n[TBF_MAX_EXTRA_HARMONICS];   /* to the built-in list     */
    float extra_harmonic_level[TBF_MAX_EXTRA_HARMONICS]; /* (src/tonegen.cpp:1602)   */
    int32_t attack_model, release_model;
    float attack_click_level, release_click_level;
    int32_t atk_clk_min_length, atk_clk_max_length; /* -1 = derive from sample rate  */
    float perc_gain_norm, perc_gain_soft, perc_gain_scaling;
    int32_t perc_bus_a, perc_bus_b, perc_bus_trig;
    /* scanner.* (defaults: reset_vibrato src/vibrato.cpp:288-305) */
    double scanner_hz;
    double scanner_amp[3];
} tbf_tonegen_spec;

typedef struct tbf_whirl_spec {
    double sample_rate;
    /* whirl.* (defaults: initValues src/whirl.cpp:40-131) */
    int32_t bypass;
    double hn_brake_pos, dr_brake_pos;
    float horn_rpm_slow, horn_rpm_fast, drum_rpm_slow, drum_rpm_fast;
    float horn_acc, horn_dec, drum_acc, drum_dec;
    double mic_angle;        /* internal value: 1 - degrees/180 (whirl.cpp:1138)     */
    float air_speed, mic_dist_cm, horn_x_offset_cm, horn_z_offset_cm;
    float horn_radius_cm, drum_radius_cm;
    int32_t lp_type;         /* drum filter: doubles in the reference                */
    double lp_f, lp_q, lp_g;
    int32_t ha_type;         /* horn filters: floats in the reference                */
    float ha_f, ha_q, ha_g;
    int32_t hb_type;
    float hb_f, hb_q, hb_g;
    float horn_level, leak_level;
    float drum_mic_width, horn_mic_width;
    int32_t rev_select;      /* whirl.speed-preset                                   */
} tbf_whirl_spec;

void tbf_tonegen_spec_default(tbf_tonegen_spec *s, double sample_rate);
void tbf_whirl_spec_default(tbf_whirl_spec *s, double sample_rate);

/* 12-TET at A4 = 440 Hz: what libMTSClient serves with no master connected
 * (libs/MTS-ESP/Client/libMTSClient.cpp:194, :879) */
void tbf_freqs_12tet(double freqs[128]);
/* Scala .scl text -> 128 note frequencies; degree 0 at MIDI 60 = 261.6255653005986 Hz
 * (what the reference's regression fixtures were generated with, tests/test_regression.py) */
int tbf_freqs_from_scl(const char *scl_text, double freqs[128]);
/* the same from the scale's n degrees as frequency ratios over degree 0 (the .scl pitch lines
 * after cents -> 2^(c/1200) and n/d -> n/d; the last one is the period) */
int tbf_freqs_from_ratios(const double *degrees, int n, double freqs[128]);
/* .scl text -> its degrees as ratios; returns n (> 0) or an error */
int tbf_scl_ratios(const char *scl_text, double *degrees, int capacity);

/* one cfg line "key=value" of the osc.*, scanner.* and whirl.* families
 * (src/tonegen.cpp:2173-2556, src/vibrato.cpp:334-357, src/whirl.cpp:992-1164).
 * Returns 1 if claimed, 0 if the key belongs to neither family, <0 on a bad value. */
int tbf_spec_apply_cfg(tbf_tonegen_spec *tg, tbf_whirl_spec *wh, const char *key,
                       const char *value);

/* Build the tables.  Arrays are owned by the tables until tbf_free_*.
 * Return TBF_ERR_REFERENCE_UB where the reference would assert/abort
 * (e.g. transformer crosstalk > 0 with 256 wheels, src/tonegen.cpp:925-937). */
int tbf_build_tonegen_tables(const tbf_tonegen_spec *spec, tbf_tonegen_tables *out);
void tbf_free_tonegen_tables(tbf_tonegen_tables *t);

/* What writeSamples() (src/tonegen.cpp:1402-1457) needs to synthesise the 256 wavetables: the
 * GPU-side init path builds everything else on the host and leaves the waves
 * -- nine tenths of initToneGenerator's time -- to tbf_gpu_tonegen_tables_create_synth(). */
typedef struct tbf_wave_synth {
    double sample_rate;
    uint32_t seed;                           /* srand() value: one rand() per sample (dither) */
    uint32_t pad_;
    double harmonics[TBF_MAX_PARTIALS];      /* wheelHarmonics after osc.harmonic.*           */
    double f1[TBF_NOF_WHEELS + 1];           /* wheel frequency, Hz                           */
    double attenuation[TBF_NOF_WHEELS + 1];  /* wheel equalisation                            */
} tbf_wave_synth;
/* tbf_build_tonegen_tables without the wave samples: out->wave_data stays NULL (offsets,
 * lengths and wave_total are filled) and *synth says how to make them; every other table,
 * the click envelopes' position in the rand() stream included, is identical. */
int tbf_build_tonegen_tables_synth(const tbf_tonegen_spec *spec, tbf_tonegen_tables *out,
                                   tbf_wave_synth *synth);
/* sample i of wheel `wheel` exactly as writeSamples computes it with the host's libm;
 * dither_bit = 1 when the sample's rand() draw was below RAND_MAX / 2 */
float tbf_wave_sample(const tbf_wave_synth *synth, int wheel, uint32_t i, int dither_bit);
/* the dither decisions of the first n draws after srand(seed), bit k of bits[k / 32] */
void tbf_dither_bits(uint32_t seed, uint64_t n, uint32_t *bits);
int tbf_build_whirl_tables(const tbf_whirl_spec *spec, tbf_whirl_tables *out);

/* setIIRFilter() (src/whirl.cpp:147-172) over eqCompute() (src/eqcomp.cpp:109-203): the
 * five float coefficients a1 a2 b0 b1 b2 of one Leslie biquad.  Returns 1 when c5 was
 * written, 0 when a parameter is outside the range the reference accepts (it then keeps
 * its previous coefficients, and so does c5), TBF_ERR_ARG for a NULL pointer. */
int tbf_iir_coefficients(float c5[5], int type, double hz, double q, double gain_db,
                         double sample_rate);
void tbf_free_whirl_tables(tbf_whirl_tables *t);

/* the extended 300-entry frequency table (src/tuning.cpp:115-147), for tests  */
int tbf_extend_frequencies(const double freqs128[128], double out[TBF_NOF_FREQS],
                           int *scale_size, float *period);
/* glibc-compatible rand() stream (TYPE_3 additive feedback), for tests        */
typedef struct tbf_rand_state {
    int32_t r[34];
    int32_t idx;
} tbf_rand_state;
void tbf_srand(tbf_rand_state *st, uint32_t seed);
int32_t tbf_rand(tbf_rand_state *st);

unsigned tbf_sizeof_tonegen_spec(void);
unsigned tbf_sizeof_whirl_spec(void);

#ifdef __cplusplus
}
#endif
#endif
