/* tbf_gpu.h -- C-ABI of the B200 batched renderer for tuneBfree's
 * tonewheel -> scanner-vibrato -> Leslie hot path.
 *
 * Plain C, plain pointers and sizes.  Every entry point cites the reference
 * interface (file:line under the reference tree) it replaces or mirrors.
 * The reference has no FFI layer of its own: its "operator API" is the set of
 * C-style engine functions over heap structs (src/tonegen.h:565-597,
 * src/vibrato.h:70-79, src/whirl.h:223-272, src/global_inst.h:32-40) that
 * the LV2 / CLAP / JACK hosts call once per 128-sample fragment
 * (src/clap.cpp:244-270, b_synth/lv2.cpp:212-239, src/main.cpp:243-292).
 *
 * All functions return 0 (TBF_OK) or a negative tbf_status; none of them
 * calls exit() (the reference exits on OOM: src/tonegen.cpp:382-386).
 */
#ifndef TBF_GPU_H
#define TBF_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TBF_FRAGMENT 128     /* BUFFER_SIZE_SAMPLES, src/tonegen.h:53           */
#define TBF_NOF_WHEELS 256   /* NOF_WHEELS, src/tonegen.h:79 (wheels are 1..256) */
#define TBF_NOF_BUSES 27     /* NOF_BUSES, src/tonegen.h:49                      */
#define TBF_NOF_DRAWBARS 9   /* NOF_DRAWBARS, src/tonegen.h:51                   */
#define TBF_MAX_KEYS 384     /* MAX_KEYS, src/tonegen.h:93                       */
#define TBF_NOF_FREQS 300    /* NOF_FREQS, src/tonegen.h:85                      */
#define TBF_MAX_PARTIALS 12  /* MAX_PARTIALS, src/tonegen.h:488                  */
#define TBF_VIB_TBL 2048     /* INCTBL_SIZE, src/vibrato.h:38                    */
#define TBF_VIB_BUF 1024     /* BUF_SIZE_BYTES, src/vibrato.h:40                 */
#define TBF_WHIRL_DISPLC 16384 /* WHIRL_DISPLC_SIZE, src/whirl.h:33              */
#define TBF_WHIRL_BUF 2048   /* WHIRL_BUF_SIZE_SAMPLES, src/whirl.h:36           */

typedef enum tbf_status {
    TBF_OK = 0,
    TBF_ERR_ARG = -1,        /* bad argument                                      */
    TBF_ERR_CUDA = -2,       /* CUDA runtime failure (tbf_gpu_last_error)         */
    TBF_ERR_NOMEM = -3,
    TBF_ERR_HANDLE = -4,     /* unknown table / instance handle                   */
    TBF_ERR_REFERENCE_UB = -5, /* input drives the reference into undefined
                                  behaviour (corePgm overflow src/tonegen.h:233,
                                  keyCompTable over-read src/tonegen.cpp:3233)     */
    TBF_ERR_GEOMETRY = -6    /* whirl geometry violates the write-ahead window the
                                kernels rely on (reference asserts: whirl.cpp:607)  */
} tbf_status;

/* ------------------------------------------------------------------------
 * Immutable per-configuration tables.  They are what initToneGenerator()
 * (src/tonegen.cpp:2905-3066) + init_vibrato() (src/vibrato.cpp:312-317)
 * leave in struct b_tonegen, flattened; many instances share one copy.
 * ---------------------------------------------------------------------- */
typedef struct tbf_tonegen_tables {
    /* wavetables: oscillators[w].wave / lengthSamples, src/tonegen.h:136-146;
     * wheel w in 1..256, index 0 unused (length 0).                          */
    uint32_t wave_offset[TBF_NOF_WHEELS + 1]; /* float offset into wave_data  */
    uint32_t wave_length[TBF_NOF_WHEELS + 1];
    const float *wave_data;
    uint64_t wave_total; /* floats in wave_data                               */

    /* keyContrib[key] linked lists (src/tonegen.h:540, built by
     * compilePlayMatrix src/tonegen.cpp:1122-1213) as CSR, list order kept.  */
    uint32_t contrib_start[TBF_MAX_KEYS + 1];
    const uint16_t *contrib_wheel;
    const uint16_t *contrib_bus;
    const float *contrib_level;

    float attack_env[TBF_NOF_DRAWBARS][TBF_FRAGMENT];  /* src/tonegen.h:246   */
    float release_env[TBF_NOF_DRAWBARS][TBF_FRAGMENT]; /* src/tonegen.h:247   */
    float key_comp_table[128];                         /* src/tonegen.h:227   */
    float draw_bar_level[TBF_NOF_BUSES][TBF_NOF_DRAWBARS]; /* src/tonegen.h:318 */

    /* percussion constants (src/tonegen.h:351-430) */
    float perc_env_gain_reset_norm, perc_env_gain_reset_soft, perc_env_scaling;
    float perc_decay_fast_norm, perc_decay_fast_soft;
    float perc_decay_slow_norm, perc_decay_slow_soft;
    float perc_drawbar_normal_gain, perc_drawbar_soft_gain;
    int32_t perc_send_bus_a, perc_send_bus_b, perc_trigger_bus;
    float output_level_trim;

    /* scanner vibrato (src/vibrato.h:42-68) */
    uint32_t vib_offset_table[3][TBF_VIB_TBL];
    uint32_t vib_stator_increment;

    /* initial run-time scalars, as left by initToneGenerator()               */
    float init_draw_bar_gain[TBF_NOF_BUSES];
    float init_swell_pedal_gain;
    float init_perc_env_gain_reset, init_perc_env_gain_decay, init_perc_drawbar_gain;
    int32_t init_perc_send_bus, init_perc_trig_restore;
    int32_t init_perc_is_soft, init_perc_is_fast, init_perc_enabled;
    uint32_t init_routing;
    int32_t init_vib_table;  /* 0..2: offset1/2/3Table; reset_vibrato -> 2     */
    int32_t init_vib_mixed;
    uint16_t init_draw_bar_change;
} tbf_tonegen_tables;

/* What initWhirl() (src/whirl.cpp:956-990) leaves in struct b_whirl
 * (src/whirl.h:65-221), flattened.                                           */
typedef struct tbf_whirl_tables {
    double sample_rate;
    const float *hn_fwd_displ, *hn_bwd_displ, *dr_fwd_displ, *dr_bwd_displ; /* 16384 each */
    const float *bfw, *bbw; /* 16384 x 5 each, struct _bw src/whirl.h:61      */
    int32_t horn_phase[6], drum_phase[6];
    float horn_spacing[6], drum_spacing[6];
    float hafw[5], hbfw[5], drfL[5], drfR[5]; /* a1 a2 b0 b1 b2               */
    /* the horn filters' type, Hz, Q, dB (haT haF haQ haG, hbT ..., src/whirl.h:173-180:
     * floats there too): what the whirl.horn.filter.* controllers start from          */
    float ha_param[4], hb_param[4];
    double mic_angle;
    float horn_acc, horn_dec, drum_acc, drum_dec;
    double rev_horn_target[9], rev_drum_target[9]; /* revoptions[]             */
    int32_t revselects[3];
    double hn_brake_pos, dr_brake_pos;
    float horn_level, leakage;
    float mic_hll, mic_hlr, mic_hrl, mic_hrr, mic_dll, mic_dlr, mic_drl, mic_drr;
    int32_t bypass;
    /* initial run-time scalars as left by initWhirl()                        */
    double init_horn_incr, init_drum_incr, init_horn_target, init_drum_target;
    double init_horn_angle, init_drum_angle;
    int32_t init_horn_acdc, init_drum_acdc, init_rev_select;
} tbf_whirl_tables;

/* ------------------------------------------------------------------------
 * Control events.  They take effect at a fragment boundary, like every host
 * of the reference does it (b_synth/lv2.cpp:1130-1136, src/main.cpp:251-266,
 * src/clap.cpp:251).
 * ---------------------------------------------------------------------- */
typedef enum tbf_event_type {
    TBF_EV_KEY_ON = 1,      /* oscKeyOn  (t, a, a)     src/tonegen.cpp:3134    */
    TBF_EV_KEY_OFF = 2,     /* oscKeyOff (t, a, a)     src/tonegen.cpp:3096    */
    TBF_EV_DRAWBAR = 3,     /* setDrawBar(t, a, b)     src/tonegen.cpp:2738    */
    TBF_EV_PERC_ENABLE = 4, /* setPercussionEnabled(a) src/tonegen.cpp:1678    */
    TBF_EV_PERC_SOFT = 5,   /* setPercussionVolume(a)  src/tonegen.cpp:1744    */
    TBF_EV_PERC_FAST = 6,   /* setPercussionFast(a)    src/tonegen.cpp:1733    */
    TBF_EV_PERC_FIRST = 7,  /* setPercussionFirst(a)   src/tonegen.cpp:1760    */
    TBF_EV_VIB_UPPER = 8,   /* setVibratoUpper(a)      src/tonegen.cpp:1635    */
    TBF_EV_VIB_LOWER = 9,   /* setVibratoLower(a)      src/tonegen.cpp:1650    */
    TBF_EV_VIB_KNOB = 10,   /* setVibratoFromInt(a)    src/vibrato.cpp:118     */
    TBF_EV_REV_OPTION = 11, /* useRevOption(w, a, 2)   src/whirl.cpp:174       */
    TBF_EV_SWELL = 12,      /* swellPedalGain = trim*a/127 src/tonegen.cpp:2885 */
    TBF_EV_REV_SELECT = 13, /* setRevSelect(w, a)      src/whirl.cpp:226       */
    /* the run-time Leslie controllers of src/whirl.cpp:975-981; a selects the rotor
     * parameter, b is the 7-bit controller value the reference's setter receives      */
    TBF_EV_WHIRL_BRAKE = 14, /* a: 0 setHornBrakePosition, 1 setDrumBrakePosition
                                (brake position = b / 127.0) src/whirl.cpp:883-893      */
    TBF_EV_WHIRL_RAMP = 15,  /* a: 0 setHornAcceleration, 1 setHornDeceleration (time
                                constant .01 + b / 80.0), 2 setDrumAcceleration,
                                3 setDrumDeceleration (.01 + b / 14.0) whirl.cpp:895-917 */
    /* one coefficient of a horn IIR as setIIRFilter() (src/whirl.cpp:147-172) leaves it:
     * a = 8 * filter (0 hafw, 1 hbfw) + index (0..4 = a1 a2 b0 b1 b2), b = the bit pattern
     * of the float.  The coefficients come out of eqCompute()'s double-precision libm
     * calls, so the host computes them (tbf_midi_compile() does, for the eight
     * whirl.horn.filter.* controllers) and the device only stores them                  */
    TBF_EV_HORN_FILTER = 16
} tbf_event_type;

typedef struct tbf_event {
    uint32_t instance; /* instance handle                                     */
    uint32_t fragment; /* fragment index inside the batch the event precedes  */
    int32_t type;      /* tbf_event_type                                      */
    int32_t a, b;
} tbf_event;

/* render flags */
#define TBF_RENDER_WHIRL_BYPASS 0x1 /* tonegen(+vibrato) only: L = R = mono buf */
#define TBF_RENDER_WHIRLPROC3 0x2   /* whirlProc3 mic-width mix (whirl.cpp:1653) */
#define TBF_RENDER_OUT_DEVICE 0x4   /* out_l/out_r are device pointers           */
#define TBF_RENDER_NO_OUTPUT 0x8    /* keep audio on device, checksums only      */
#define TBF_RENDER_WHIRL_INPUT 0x10 /* Leslie only: whirlProc / whirlProc3 on the caller's mono
                                       buffers `in` (src/whirl.h:232-249) -- the second half of
                                       the hosts' chain oscGenerateFragment -> preamp -> reverb
                                       -> whirlProc3 (src/clap.cpp:254-258) when the reverb or
                                       the overdrive stays on the CPU.  The tone generator's
                                       state is not touched; only rotor events are applied      */
#define TBF_RENDER_STEMS 0x20       /* whirlProc2's separate horn / drum outputs
                                       (src/whirl.h:232-238, src/whirl.cpp:1598-1612): outHL, outHR
                                       = hornLevel * horn line + leak, outDL, outDR = the filtered
                                       drum lines, written to stems[0..3] next to out_l / out_r   */
#define TBF_RENDER_OUT_CHUNKED 0x40 /* host output in chunk-major layout: the batch's fragments are
                                       cut into chunks of tbf_gpu_get_host_chunk() fragments and
                                       chunk c of instance i lies at out_x + (c * n_instances + i) *
                                       chunk_samples -- every chunk then crosses the host link as ONE
                                       contiguous copy per channel instead of n_instances strided rows
                                       (out_stride is ignored; the last chunk may be shorter and keeps
                                       the row length chunk_samples)                                */

/* Testing and measurement only: variants of stage T that all produce the same bits.  0x100: every
 * fragment takes the full active-list pass; 0x200: quiet fragments one at a time, not in pairs;
 * 0x400: every bus sum of a quiet fragment is taken again over all records; 0x800: quiet fragments
 * by the one merged loop, never bus by bus; 0x4000: quiet fragments sum all three buses, read or not;
 * bits 12-13: how the wheels' samples are loaded (0 = the library's choice, 1 = TMA bulk copies,
 * 2 = per-lane cp.async, 3 = direct loads); 0x8000: stages R and D move their 32 x 32 tiles with
 * cp.async and store loops instead of 2-D TMA boxes. */
#define TBF_RENDER_DEBUG_FULL_PASS 0x100
#define TBF_RENDER_DEBUG_NO_PAIRS 0x200
#define TBF_RENDER_DEBUG_BUS_ALL 0x400
#define TBF_RENDER_DEBUG_NO_BUS 0x800
#define TBF_RENDER_DEBUG_T_TMA 0x1000
#define TBF_RENDER_DEBUG_T_CPASYNC 0x2000
#define TBF_RENDER_DEBUG_T_DIRECT 0x3000
#define TBF_RENDER_DEBUG_ALL_BUSES 0x4000
#define TBF_RENDER_DEBUG_NO_TMA_TILES 0x8000

typedef struct tbf_batch_desc {
    uint32_t n_instances;
    const uint32_t *instance_ids; /* NULL = instances 0..n_instances-1         */
    uint32_t n_fragments;         /* 128-sample fragments to render per instance */
    const tbf_event *events;      /* sorted by (position in instance_ids, fragment);
                                     order inside one (instance, fragment) is kept */
    uint64_t n_events;
    float *out_l, *out_r; /* [n_instances][out_stride] floats                  */
    uint64_t out_stride;  /* floats between instances, >= n_fragments*128       */
    uint32_t flags;
    uint64_t *checksums;  /* optional, host, [n_instances]: FNV-style hash of the
                             rendered L/R bit patterns                          */
    const float *in;      /* TBF_RENDER_WHIRL_INPUT: [n_instances][in_stride] mono input,
                             host memory, or device memory with TBF_RENDER_OUT_DEVICE   */
    uint64_t in_stride;   /* floats between instances, >= n_fragments*128       */
    float *stems[4];      /* TBF_RENDER_STEMS: outHL outHR outDL outDR, same layout and memory
                             space as out_l / out_r; NULL entries are skipped      */
} tbf_batch_desc;

typedef struct tbf_ctx tbf_ctx;

/* context = one GPU (one host thread / process per GPU; not re-entrant)      */
int tbf_gpu_create(tbf_ctx **out, int device);
void tbf_gpu_destroy(tbf_ctx *ctx);
const char *tbf_gpu_last_error(const tbf_ctx *ctx);

/* upload table sets; the library copies everything                          */
int tbf_gpu_tonegen_tables_create(tbf_ctx *ctx, const tbf_tonegen_tables *t, int *handle);
int tbf_gpu_whirl_tables_create(tbf_ctx *ctx, const tbf_whirl_tables *w, int *handle);
/* GPU-side init: the same as tbf_gpu_tonegen_tables_create, but the 256
 * wavetables -- writeSamples(), src/tonegen.cpp:1402-1457, nine tenths of initToneGenerator's
 * time -- are synthesised on the device from `synth` (include/tbf_tables.h:
 * tbf_build_tonegen_tables_synth leaves t->wave_data NULL and fills it).  Bit-identical to the
 * host-built tables: every double operation is the reference's; samples whose float could
 * hinge on the last bit of a libm sine (about one in ten table sets) are recomputed by the
 * host's libm and patched.  tbf_gpu_synth_stats: how many samples were flagged / differed. */
struct tbf_wave_synth;
int tbf_gpu_tonegen_tables_create_synth(tbf_ctx *ctx, const tbf_tonegen_tables *t,
                                        const struct tbf_wave_synth *synth, int *handle);
int tbf_gpu_synth_stats(const tbf_ctx *ctx, uint64_t *flagged, uint64_t *differed);
/* initToneGenerator for MANY organs at once (BASELINE config 4: one tuning per organ, the
 * reference re-runs initToneGenerator per tuning: src/clap.cpp:129-160, src/tonegen.cpp:2905):
 * n_threads host threads (<= 0: one per core) build the table sets of specs[0..n) while the
 * calling thread uploads them in order and creates one instance each with Leslie set
 * `whirl_tables`; device_waves != 0 synthesises the wavetables on the GPU.  tonegen_handles
 * (may be NULL) and instance_ids receive n entries; stats (may be NULL) = {wall seconds, seconds
 * the uploader waited for builders, seconds it spent uploading}.  On an error the organs created
 * so far stay; the code is the first failing call's.  (tbf_pipeline.cpp) */
struct tbf_tonegen_spec;
int tbf_gpu_organs_create(tbf_ctx *ctx, const struct tbf_tonegen_spec *specs, uint32_t n, int whirl_tables,
                          int n_threads, int device_waves, int *tonegen_handles, uint32_t *instance_ids,
                          double *stats);
/* the raw waves of a set as the device holds them (wave_offset layout): parity tests */
int tbf_gpu_tonegen_tables_read_waves(tbf_ctx *ctx, int handle, float *out, uint64_t n_floats);
/* how the kernels treat a whirl table set: 1 = parallel ordered gather of the delay lines
 * (the three taps of each line have disjoint delay ranges), 0 = serial path; ring = shared-
 * memory slots per delay line (512 or 2048); window = live slots ahead of the reader */
int tbf_gpu_whirl_tables_info(tbf_ctx *ctx, int handle, int *gather_ok, int *ring_size, int *window);

/* replaces allocTonegen+initToneGenerator+allocWhirl+initWhirl as far as the
 * run-time state goes (src/tonegen.cpp:3780, :2905; src/whirl.cpp:136, :956) */
int tbf_gpu_instance_create(tbf_ctx *ctx, int tonegen_tables, int whirl_tables, uint32_t count,
                            uint32_t *first_id);
int tbf_gpu_instance_count(const tbf_ctx *ctx);
/* freeToneGenerator + freeWhirl (src/tonegen.cpp:3079-3090, src/whirl.cpp:140-145) as far as the
 * run-time state goes: the slots are recycled by later tbf_gpu_instance_create calls, and the
 * instances stop counting as users of their table sets                                        */
int tbf_gpu_instance_destroy(tbf_ctx *ctx, uint32_t first_id, uint32_t count);
/* release a table set.  Table sets are reference-counted by their instances: the device memory
 * is freed (and the handle recycled) once the set has been destroyed AND its last instance is
 * gone, whichever comes last -- the order a host calls freeToneGenerator / freeWhirl in does not
 * matter.  The CLAP host re-creates the tone generator on every tuning change
 * (src/clap.cpp:129-160): destroy + create keeps HBM flat                                     */
int tbf_gpu_tonegen_tables_destroy(tbf_ctx *ctx, int handle);
int tbf_gpu_whirl_tables_destroy(tbf_ctx *ctx, int handle);
/* live objects and device memory of the context's tables and instance array (lifetime tests) */
int tbf_gpu_resource_counts(const tbf_ctx *ctx, uint32_t *live_instances, uint32_t *instance_slots,
                            uint32_t *live_tonegen_sets, uint32_t *live_whirl_sets);
/* workload statistics of instances (NULL ids = 0..n-1): fragments rendered since create / reset
 * and the sum over them of the active-list length (oscGenerateFragment's activeOscLEnd after
 * the message-queue drain, src/tonegen.cpp:3343); summed over the listed instances            */
int tbf_gpu_instance_stats(tbf_ctx *ctx, const uint32_t *ids, uint32_t n, uint64_t *fragments,
                           uint64_t *active_wheels);
/* back to the state right after create (what a host gets by re-creating the
 * engine, b_synth/lv2.cpp:1080-1093)                                         */
int tbf_gpu_instance_reset(tbf_ctx *ctx, uint32_t first_id, uint32_t count);

/* the batched replacement for the per-fragment chain
 * oscGenerateFragment -> [vibratoProc] -> whirlProc/whirlProc3
 * (src/clap.cpp:254-258)                                                     */
int tbf_gpu_render_batch(tbf_ctx *ctx, const tbf_batch_desc *desc);

/* NO COUNTERPART IN THE REFERENCE: added for callers that
 * may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  tbf_gpu_render_batch's chain for one SMALL batch with every intermediate buffer
 * (stage streams, per-fragment records, the uploaded events, ids, checksums) in one file-scope
 * static device array of tbf_gpu_noalloc_scratch_bytes() bytes (6.4 MB; every GPU that loads
 * the library holds it until exit): the call makes no cudaMalloc / cudaFree and calls no
 * allocating routine.  Limits: n_instances <= TBF_NOALLOC_MAX_INSTANCES, n_fragments <=
 * TBF_NOALLOC_MAX_FRAGMENTS, n_events <= TBF_NOALLOC_MAX_EVENTS per call -- larger batches are
 * rejected with TBF_ERR_ARG and a message naming the limits.  flags: TBF_RENDER_OUT_DEVICE is
 * required (out_l / out_r are the caller's device buffers), plus WHIRL_BYPASS / WHIRLPROC3;
 * checksums (host) are supported; `in` and `stems` are not.  The instances keep their state
 * from call to call exactly as with tbf_gpu_render_batch (results are bit-identical to it).
 * The array is shared by every call and every context on the device and keeps the last call's
 * contents; what must start at zero is cleared per call.  Calls that overlap on the GPU race:
 * the function runs on the context's stream and returns after that stream has drained; do not
 * call it from two threads for the same device at once.                                      */
#define TBF_NOALLOC_MAX_INSTANCES 32
#define TBF_NOALLOC_MAX_FRAGMENTS 32
#define TBF_NOALLOC_MAX_EVENTS 4096
int tbf_gpu_render_small_noalloc(tbf_ctx *ctx, const tbf_batch_desc *desc);
unsigned long long tbf_gpu_noalloc_scratch_bytes(void);

/* asynchronous pieces: prepare validates the batch and uploads events / ids once;
 * launch renders the prepared batch on the library's stream and may be repeated
 * (the instances keep playing the same event script); render_batch_launch =
 * prepare + launch; render_batch = prepare + launch + synchronize.
 * Host outputs are streamed back in chunks overlapping the render (page-locked
 * buffers from tbf_gpu_host_alloc make that copy asynchronous).               */
int tbf_gpu_batch_prepare(tbf_ctx *ctx, const tbf_batch_desc *desc);
int tbf_gpu_batch_launch(tbf_ctx *ctx);
int tbf_gpu_render_batch_launch(tbf_ctx *ctx, const tbf_batch_desc *desc);
int tbf_gpu_synchronize(tbf_ctx *ctx);
void *tbf_gpu_host_alloc(tbf_ctx *ctx, size_t bytes);
void tbf_gpu_host_free(tbf_ctx *ctx, void *p);
/* fragments per device -> host staging chunk: get returns what TBF_RENDER_OUT_CHUNKED cuts the
 * caller's buffers by (32 unless set); set applies to both host layouts (the row layout's default is
 * 125) */
int tbf_gpu_set_host_chunk(tbf_ctx *ctx, uint32_t fragments);
uint32_t tbf_gpu_get_host_chunk(const tbf_ctx *ctx);
/* page-locked host buffer options (tbf_gpu_host_alloc2): write-combined pages are faster for the
 * device to fill and slow for the host to read back -- for hosts that hand the audio on (file,
 * socket) without touching it                                                                  */
#define TBF_HOST_WRITE_COMBINED 0x1
void *tbf_gpu_host_alloc2(tbf_ctx *ctx, size_t bytes, unsigned flags);
/* What tbf_gpu_batch_prepare does with a batch's event list before anything reaches the device, for
 * callers that want to check a list ahead of time (needs no context and no GPU): the events must be
 * sorted by (slot, fragment) with slot < n_instances and fragment < n_fragments; start[s] receives
 * the index of the first event of a slot >= s, start[n_instances] = n_events (n_instances + 1
 * entries).  TBF_OK, or TBF_ERR_ARG with *why (may be null) naming the rule that was broken.      */
int tbf_gpu_event_index(const tbf_event *events, uint64_t n_events, uint32_t n_instances,
                        uint32_t n_fragments, uint64_t *start, const char **why);
/* device -> host link ceiling of this context's GPU: plain 1-D page-locked cudaMemcpyAsync of
 * `bytes`, `reps` timed passes after one untimed, the mean rate in GB/s                        */
int tbf_gpu_measure_d2h(tbf_ctx *ctx, size_t bytes, int reps, double *gbs);
/* CUDA-event stopwatch on the library's stream                                 */
int tbf_gpu_timer_start(tbf_ctx *ctx);
int tbf_gpu_timer_stop(tbf_ctx *ctx, float *ms);
/* device time of the last launch in ms (CUDA events on the library's stream),
 * and how many kernels it launched                                           */
int tbf_gpu_last_launch_ms(tbf_ctx *ctx, float *ms, int *n_kernel_launches);

/* per-stage device times: while profiling is on every stage kernel (T tone generator +
 * scanner, R ramps + horn filters + rotors, W Leslie taps + delay lines, D drum filters +
 * mix) runs alone between two CUDA events; times accumulate until profiling is reset  */
int tbf_gpu_set_stage_profiling(tbf_ctx *ctx, int on);
int tbf_gpu_stage_times(tbf_ctx *ctx, float ms[4], int launches[4]);

/* parity debugging: integer + float run-time state of one instance           */
typedef struct tbf_state_export {
    int32_t active_osc_l_end;
    int32_t active_osc_list[TBF_NOF_WHEELS + 1];
    uint32_t osc_pos[TBF_NOF_WHEELS + 1];
    int32_t osc_acl_pos[TBF_NOF_WHEELS + 1];
    int32_t ref_count[TBF_NOF_WHEELS + 1];
    uint32_t old_routing, new_routing, upper_key_count;
    int32_t key_down_count;
    float perc_env_gain, key_comp_level, pz;
    uint32_t vib_stator, vib_out_pos;
    float vib_buffer[TBF_VIB_BUF];
    uint32_t whirl_outpos;
    double horn_angle, drum_angle, horn_incr, drum_incr;
    int32_t horn_acdc, drum_acdc;
    float whirl_buf[4][TBF_WHIRL_BUF]; /* HL HR DL DR                          */
    float iir_z[4][2];                 /* hafw hbfw drfL drfR: z0 z1           */
    float horn_iir[2][5];              /* hafw hbfw: a1 a2 b0 b1 b2 (per instance:
                                          TBF_EV_HORN_FILTER)                   */
    double brake_pos[2];               /* hnBrakePos drBrakePos                */
    double ramp_one_minus_l[4];        /* 1 - exp(-1 / (SR / 128 * tau)), tau = hornAcc
                                          hornDec drumAcc drumDec               */
    float filt_z[4];
    float adx[3][8];
    int32_t adi[3];
    uint32_t error_flags;              /* TBF_IFLAG_*                          */
    uint32_t pad_;
    uint64_t stat_fragments;           /* fragments rendered since create/reset */
    uint64_t stat_active_wheels;       /* sum over them of the active-list length */
} tbf_state_export;
int tbf_gpu_export_state(tbf_ctx *ctx, uint32_t instance, tbf_state_export *out);

#define TBF_IFLAG_COREPGM_OVERFLOW 0x1 /* > CR_PGMMAX core instrs: reference UB */
#define TBF_IFLAG_KEYCOMP_OOB 0x2      /* keyDownCount >= 128: reference over-read */
#define TBF_IFLAG_BAD_EVENT 0x4
#define TBF_IFLAG_GATHER_FALLBACK 0x8  /* informational: some fragment used the serial
                                          delay-line path (rotor faster than the gather covers) */

#ifdef __cplusplus
}
#endif
#endif /* TBF_GPU_H */
